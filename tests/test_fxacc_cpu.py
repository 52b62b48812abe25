"""csrc/fxacc.cuh on the host (tests/fxacc_emulation.py): sums held in 32-bit limbs are EXACT -- for values inside the
window the result equals math.fsum of the group bit for bit, whatever the order and however much cancels; integer sums
wrap like Spark's long; values outside the window are reported to the caller and never forced in."""
import math

import numpy as np
import pytest

import fxacc_emulation


@pytest.fixture(scope="module")
def emul(tmp_path_factory):
    return fxacc_emulation.build(str(tmp_path_factory.mktemp("fx_emul")))


def _fsum_by_group(v, g, ngroups):
    return np.array([math.fsum(v[g == k].tolist()) for k in range(ngroups)])


@pytest.mark.parametrize("scale,seed", [(1.0, 1), (1e3, 2), (1e-200, 3), (1e200, 4), (2.0**-1060, 5)])
def test_in_window_sums_equal_fsum_bit_for_bit(emul, scale, seed):
    rng = np.random.default_rng(seed)
    n, ngroups = 200_000, 37
    # 55 binades of spread below the maximum (the window keeps 62 below the sampled maximum), both signs, heavy cancellation in some groups
    v = rng.choice([-1.0, 1.0], n) * rng.uniform(0.5, 1.0, n) * scale * np.exp2(-rng.integers(0, 55, n).astype(np.float64))
    g = rng.integers(0, ngroups, n).astype(np.int32)
    v[g == 3] = np.concatenate([x := v[g == 3][: (g == 3).sum() // 2], -x, np.zeros((g == 3).sum() - 2 * len(x))])   # sums to exactly 0
    v[g == 4] = np.abs(v[g == 4])
    out, spill, nspilled, _ = fxacc_emulation.group_sums(emul, v, g, ngroups, np.abs(v).max())
    assert nspilled == 0 and not spill.any()
    want = _fsum_by_group(v, g, ngroups)
    assert np.array_equal(out.view(np.uint64), want.view(np.uint64))
    assert out[3] == 0.0
    # the order in which the adds land does not matter
    perm = rng.permutation(n)
    out2, _, _, _ = fxacc_emulation.group_sums(emul, v[perm], g[perm], ngroups, np.abs(v).max())
    assert np.array_equal(out2.view(np.uint64), out.view(np.uint64))


def test_values_outside_the_window_are_handed_back(emul):
    rng = np.random.default_rng(9)
    n, ngroups = 50_000, 5
    v = rng.standard_normal(n) * 1e3
    g = rng.integers(0, ngroups, n).astype(np.int32)
    smax = np.abs(v).max()                                   # what the sample saw
    special = {10: np.nan, 11: np.inf, 12: -np.inf, 13: smax * 2.0**13, 14: smax * 2.0**12, 15: 1e-30, 16: 5e-324, 17: -0.0, 18: 0.0}
    for i, x in special.items():
        v[i] = x
    out, spill, nspilled, e_base = fxacc_emulation.group_sums(emul, v, g, ngroups, smax)
    # NaN, +-Inf, the value 13 binades above the sampled maximum (headroom is 12) and the two far below the window are
    # not placed; the value 12 binades above, and the zeros, are
    assert nspilled == 6
    inside = np.ones(n, bool)
    inside[[10, 11, 12, 13, 15, 16]] = False
    want = _fsum_by_group(v[inside], g[inside], ngroups)
    assert np.array_equal(out.view(np.uint64), want.view(np.uint64))
    assert 1 <= e_base <= 2046 - 74


def test_window_width_and_row_count_headroom(emul):
    # 74 binades are held exactly: the smallest and the largest placeable magnitude in one group, 2^20 times each
    smax = 2.0**40
    _, _, _, e_base = fxacc_emulation.group_sums(emul, np.array([1.0]), np.array([0], np.int32), 1, smax)
    top, bottom = 2.0**(e_base - 1023 + 74), 2.0**(e_base - 1023)
    n = 1 << 20
    v = np.concatenate([np.full(n, top * (2 - 2.0**-52)), np.full(n, bottom * (1 + 2.0**-52))])
    out, _, nspilled, _ = fxacc_emulation.group_sums(emul, v, np.zeros(2 * n, np.int32), 1, smax)
    assert nspilled == 0 and out[0] == math.fsum(v.tolist())
    out, _, nspilled, _ = fxacc_emulation.group_sums(emul, np.array([top * 2, bottom / 2]), np.zeros(2, np.int32), 1, smax)
    assert nspilled == 2 and out[0] == 0.0


def test_integer_sums_wrap_like_a_jvm_long_and_round_once_as_double(emul):
    rng = np.random.default_rng(4)
    n, ngroups = 100_000, 11
    v = rng.integers(-2**63, 2**63 - 1, n, dtype=np.int64)
    v[:5] = [2**63 - 1, 2**63 - 1, -2**63, -1, 0]
    g = rng.integers(0, ngroups, n).astype(np.int32)
    g[:5] = 0
    low, dbl = fxacc_emulation.int_sums(emul, v, g, ngroups)
    for k in range(ngroups):
        exact = sum(int(x) for x in v[g == k])
        assert int(low[k]) == ((exact + 2**63) % 2**64) - 2**63          # Spark's sum(bigint) wraps
        assert dbl[k] == float(exact)                                    # the 96-bit value rounded once
