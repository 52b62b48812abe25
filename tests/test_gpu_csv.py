"""GPU CSV decode (fsx_csv_decode) against the host reader (host/csv.hpp): inferred types and every cell
bit for bit, on the tour's schemas and on adversarial numeric text (rounding boundaries, subnormals, 17-19 digit
mantissas, exponents, signs, nulls, CRLF, blank lines)."""
import os
import random
import struct

import numpy as np
import pytest

import fsx
from fsx import tour

pytestmark = pytest.mark.gpu


def _write(tmp_path, header, rows, eol="\n", trailing=True):
    p = tmp_path / "t.csv"
    body = eol.join(rows) + (eol if trailing else "")
    p.write_bytes((header + eol + body).encode())
    return str(p), body.encode(), header.count(",") + 1


def _compare(ctx, path, body, ncols):
    types, t = ctx.csv_decode(body, ncols)
    for c in range(ncols):
        hty, cells, valid = tour.read_csv_column(path, c)
        assert types[c] == hty, f"column {c}: inferred {types[c]}, host {hty}"
        data, gvalid = t.to_numpy(c, True)
        assert t.nrows == len(cells) and np.array_equal(gvalid, valid), f"column {c}: nulls"
        if hty == "string":
            continue
        bits = data.view(np.uint64) if data.dtype.itemsize == 8 else data.astype(np.int64).view(np.uint64)
        bad = np.nonzero((bits != cells) & valid)[0]
        assert len(bad) == 0, f"column {c} row {bad[:3]}: gpu {bits[bad[:3]]} host {cells[bad[:3]]}"
    return types, t


def test_tour_shaped_tables_decode_like_the_host_reader(ctx, tmp_path):
    rng = random.Random(7)
    players = [f"{rng.randint(18, 38)},{rng.lognormvariate(4, 0.3)!r},{rng.randint(1, 50)},{rng.lognormvariate(13, 1)!r}" for _ in range(1250)]
    path, body, nc = _write(tmp_path, "age,rating,team_id,worth", players)
    types, _ = _compare(ctx, path, body, nc)
    assert types == ["int", "double", "int", "double"]
    teams = [f"{rng.uniform(1e6, 1e8):.2f},{i + 1},team {i},owner_{i % 7},{50 - i}" for i in range(50)]
    path, body, nc = _write(tmp_path, "team_budget,team_id,team_name,team_owner,team_position", teams, eol="\r\n")
    types, _ = _compare(ctx, path, body, nc)
    assert types == ["double", "int", "string", "string", "int"]


def test_adversarial_numeric_text_is_bit_identical_to_strtod(ctx, tmp_path):
    rng = random.Random(11)
    specials = ["9007199254740993", "9007199254740995", "1e23", "8.98846567431158e307", "4.9e-324", "2.4703282292062327e-324",
                "2.4703282292062328e-324", "1.7976931348623157e308", "1.7976931348623159e308", "0.1", "2.2250738585072011e-308",
                "2.2250738585072014e-308", "-0", "-0.0", "+5", "5.", ".5", "1E+5", "1e-400", "1e400", "NaN", "Infinity", "-Infinity", "", "000123.4500",
                "0.000000000000000000000000000001", "123456789012345678", "1234567890123456789", "0.30000000000000004", "5e-324", "3e-324"]
    rows = []
    for i in range(60_000):
        nd = rng.randint(1, 19)
        digs = "".join(rng.choice("0123456789") for _ in range(nd))
        k = rng.randint(0, nd)
        s = digs[:k] + "." + digs[k:] if rng.random() < 0.7 else digs
        if s == ".":
            s = "0"
        if rng.random() < 0.5:
            s += f"e{rng.randint(-330, 310)}"
        if rng.random() < 0.3:
            s = "-" + s
        d = specials[i % len(specials)] if i % 37 == 0 else s
        big = str(rng.randint(-2**63, 2**63 - 1)) if i % 5 else ""
        small = str(rng.randint(-2**31, 2**31 - 1)) if i % 11 else ""
        mixed = str(rng.randint(-10, 10)) if i % 3 else f"{rng.random()}"
        rows.append(f"{d},{big},{small},{mixed}")
        if i % 5000 == 0:
            rows.append("")                                   # blank lines are skipped
    path, body, nc = _write(tmp_path, "d,l,i,m", rows, trailing=False)
    types, _ = _compare(ctx, path, body, nc)
    assert types == ["double", "bigint", "int", "double"]
    # widening rules: one long cell makes an int column bigint; one non-number makes it string; overflow beyond int64 -> double
    path, body, nc = _write(tmp_path, "a,b,c", ["1,2,3", "3000000000,x,99999999999999999999", ",,"])
    types, t = _compare(ctx, path, body, nc)
    assert types == ["bigint", "string", "double"] and t.nrows == 3


def test_files_the_decoder_declines_and_ragged_rows(ctx):
    with pytest.raises(fsx.FsxError) as e:
        ctx.csv_decode(b'1,"a,b"\n', 2)
    assert e.value.code == fsx.api.E_UNSUPPORTED
    with pytest.raises(fsx.FsxError) as e:
        ctx.csv_decode(b"1,2\n3\n", 2)
    assert e.value.code == fsx.api.E_DATA and "row 2" in e.value.message
    types, t = ctx.csv_decode(b"", 3)
    assert t.nrows == 0 and types == ["string"] * 3
    # 25 digits on a rounding boundary cannot be decided from 19 of them: declined, not guessed
    with pytest.raises(fsx.FsxError):
        ctx.csv_decode(b"9007199254740993.0000000001\n", 1)
    types, t = ctx.csv_decode(b"1234567890123456789012345\n", 1)       # 25 digits well inside one rounding interval
    assert types == ["double"] and t.to_numpy(0)[0] == float("1234567890123456789012345")
