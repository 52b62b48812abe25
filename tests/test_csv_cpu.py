"""The cell grammar of the GPU CSV decoder (csrc/csv.cu: parse_i64, scan_decimal, Eisel-Lemire, parse_f64, cell_class)
compiled for the host (tests/csv_emulation.py) against Python's int() and correctly rounded float(): the same bits as
strtod / Java's Double.parseDouble, which is what Spark's CSV scan before ComputeFeatures.scala:104 returns.  A cell
the conversion declines (`decided` false: it cannot prove the rounding) must be rare and is the only allowed difference."""
import random
import re
import struct
from fractions import Fraction

import numpy as np
import pytest

import csv_emulation

CLS = {"null": 0, "int": 1, "long": 2, "double": 3, "string": 4}
JAVA_DOUBLE = re.compile(r"[+-]?(\d+(\.\d*)?|\.\d+)([eE][+-]?\d+)?\Z", re.ASCII)       # the decimal forms of Double.parseDouble (no suffix, no hex)


@pytest.fixture(scope="module")
def emul(tmp_path_factory):
    return csv_emulation.build(str(tmp_path_factory.mktemp("csv_emul")))


def _bits(x: float) -> int:
    return struct.unpack("<Q", struct.pack("<d", x))[0]


def _check_doubles(emul, cells, max_undecided=0):
    r = csv_emulation.parse_cells(emul, cells)
    want = np.array([_bits(float(c)) for c in cells], np.uint64)
    assert r["f_ok"].all()
    decided = r["f_decided"].astype(bool)
    bad = np.nonzero(decided & (r["f_bits"] != want))[0]
    assert len(bad) == 0, [(cells[i], hex(int(r["f_bits"][i])), hex(int(want[i]))) for i in bad[:5]]
    assert int((~decided).sum()) <= max_undecided, [cells[i] for i in np.nonzero(~decided)[0][:5]]
    return r


def test_known_rounding_cases_and_limits(emul):
    cells = ["9007199254740993", "9007199254740995", "1e23", "8.98846567431158e307", "4.9e-324", "2.4703282292062327e-324",
             "2.4703282292062328e-324", "1.7976931348623157e308", "1.7976931348623158e308", "1.7976931348623159e308", "0.1", "2.2250738585072011e-308",
             "2.2250738585072014e-308", "-0", "-0.0", "+5", "5.", ".5", "1E+5", "1e-400", "1e400", "000123.4500", "0.000000000000000000000000000001",
             "123456789012345678", "1234567890123456789", "0.30000000000000004", "5e-324", "3e-324", "2e-324", "1e-323", "6.631236871469758e-316",
             "3.237883913302901e-319", "1448997445238699", "17976931348623157e292", "0e0", "0.0e999", "1e22", "1e-22", "9.5e-1", "7.2057594037927933e16",
             "9223372036854775807", "9223372036854775808", "18446744073709551615", "18446744073709551616", "1" + "0" * 30, "0." + "0" * 30 + "1"]
    r = _check_doubles(emul, cells)
    assert not r["f_bits"][cells.index("1e-400")] and r["f_bits"][cells.index("1e400")] == 0x7FF0000000000000
    assert r["f_bits"][cells.index("-0")] == 0x8000000000000000


def test_special_words_and_what_is_not_a_number(emul):
    cells = ["NaN", "Infinity", "+Infinity", "-Infinity", "nan", "inf", "-inf", "0x10", "1e", "e5", ".", "+", "-", "1.2.3", "1e+", " 1", "1 ", "1,5", "--1", "1d",
             "1f", "١٢", "1_000", "", "abc"]
    r = csv_emulation.parse_cells(emul, cells)
    ok = dict(zip(cells, r["f_ok"].astype(bool)))
    assert [c for c in cells if ok[c]] == ["NaN", "Infinity", "+Infinity", "-Infinity"]
    assert r["f_bits"][0] == 0x7FF8000000000000 and r["f_bits"][1] == r["f_bits"][2] == 0x7FF0000000000000 and r["f_bits"][3] == 0xFFF0000000000000
    cls = dict(zip(cells, r["cls"]))
    assert cls[""] == CLS["null"] and cls["NaN"] == CLS["double"] and cls["abc"] == cls["1d"] == cls["0x10"] == cls[" 1"] == CLS["string"]


def test_integers_follow_java_parse_long(emul):
    rng = random.Random(3)
    cells = ["0", "-0", "+0", "007", "-2147483648", "2147483647", "2147483648", "-2147483649", "9223372036854775807", "-9223372036854775808",
             "9223372036854775808", "-9223372036854775809", "+", "-", "1-", "12a", "1.0", "99999999999999999999", "+42", "٣"]
    cells += [str(rng.randint(-2**63, 2**63 - 1)) for _ in range(20_000)] + [str(rng.randint(-2**33, 2**33)) for _ in range(20_000)]
    cells += [("+" if rng.random() < 0.5 else "") + "0" * rng.randint(0, 5) + str(rng.randint(0, 2**64)) for _ in range(5_000)]
    r = csv_emulation.parse_cells(emul, cells)
    for i, c in enumerate(cells):
        good = re.fullmatch(r"[+-]?[0-9]+", c, re.ASCII) is not None and -2**63 <= int(c) < 2**63
        assert bool(r["i_ok"][i]) == good, c
        if good:
            v = int(c)
            assert int(r["i_val"][i]) == v, c
            assert r["cls"][i] == (CLS["int"] if -2**31 <= v < 2**31 else CLS["long"]), c
        elif JAVA_DOUBLE.match(c):
            assert r["cls"][i] == CLS["double"], c
        else:
            assert r["cls"][i] == CLS["string"], c


def test_random_literals_round_like_strtod(emul):
    rng = random.Random(11)
    cells = []
    for _ in range(300_000):
        nd = rng.randint(1, 19)
        digs = "".join(rng.choice("0123456789") for _ in range(nd))
        point = rng.randint(0, nd)
        body = digs[:point] + ("." + digs[point:] if rng.random() < 0.7 or point == 0 else digs[point:])
        if body.startswith("."):
            body = ("0" if rng.random() < 0.5 else "") + body
        exp = rng.choice(["", "", f"e{rng.randint(-330, 310)}", f"E{rng.randint(-30, 30):+d}"])
        cells.append(rng.choice(["", "", "-", "+"]) + body + exp)
    cells += [repr(struct.unpack("<d", struct.pack("<Q", rng.getrandbits(64) & 0x7FEFFFFFFFFFFFFF))[0]) for _ in range(200_000)]     # shortest round-trip forms
    cells += [f"{rng.lognormvariate(0, 20):.17e}" for _ in range(100_000)]
    _check_doubles(emul, cells)


def test_values_next_to_rounding_boundaries(emul):
    """the decimal expansion of the exact midpoint between two neighbouring doubles, cut or nudged at 17-19 and at 25-40 digits:
    at most 19 digits are always decided; longer mantissas may be declined only when the cut straddles the boundary"""
    rng = random.Random(5)
    short, long_ = [], []
    for _ in range(20_000):
        bits = rng.getrandbits(64) & 0x7FEFFFFFFFFFFFFF
        if rng.random() < 0.15:
            bits &= 0x001FFFFFFFFFFFFF                                # subnormals and the smallest normals
        lo = struct.unpack("<d", struct.pack("<Q", bits))[0]
        hi = struct.unpack("<d", struct.pack("<Q", bits + 1))[0]
        mid = (Fraction(lo) + Fraction(hi)) / 2
        # scientific expansion of the midpoint to 45 significant digits
        e10 = len(str(mid.numerator)) - len(str(mid.denominator))
        scaled = mid / Fraction(10) ** e10
        if scaled < 1:
            e10 -= 1
            scaled *= 10
        digits = str(int(scaled * 10**44))
        for nd in (17, 18, 19):
            d = int(digits[:nd]) + rng.choice([0, 0, 1, -1])
            short.append(f"{d}e{e10 - nd + 1}")
        nd = rng.randint(20, 40)
        tail = rng.choice(["", "0", "1", "9", "5", "49", "51"])
        long_.append(f"{digits[0]}.{digits[1:nd]}{tail}e{e10}")
    _check_doubles(emul, short)
    r = _check_doubles(emul, long_, max_undecided=len(long_))            # wrong bits are never allowed; declining is
    # these cells sit ON a boundary by construction, so many are declined.  A random 25-digit mantissa is declined only when
    # its first 19 digits w and w + 1 round differently: boundaries are ~2^10 units of w apart, so about one cell in a thousand
    far = [f"{rng.randint(10**24, 10**25 - 1)}e{rng.randint(-300, 280)}" for _ in range(50_000)]
    r = _check_doubles(emul, far, max_undecided=len(far) // 250)
    assert r["f_ok"].all()
