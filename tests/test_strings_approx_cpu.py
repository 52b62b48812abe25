"""Host-side helpers of SURVEY 8(f)-1 that need no GPU: dictionary encoding of string columns and the HyperLogLog
estimator, both inside libfsx.so, against Python / the oracle."""
import numpy as np

import fsx
from oracle import oracle as orc


def test_dictionary_encode_first_appearance_order_nulls_and_unicode():
    strings = ["Residential", "Condo", None, "Residential", "", "Multi-Family", "Condo", "Ünïcode", None]
    codes, dictionary = fsx.dictionary_encode(strings)
    assert dictionary == ["Residential", "Condo", "", "Multi-Family", "Ünïcode"]
    assert codes.tolist() == [0, 1, -1, 0, 2, 3, 1, 4, -1]
    assert [dictionary[c] if c >= 0 else None for c in codes] == strings
    codes, dictionary = fsx.dictionary_encode([])
    assert len(codes) == 0 and dictionary == []
    rng = np.random.default_rng(1)
    pool = [f"team_{i}" for i in range(500)]
    big = [pool[i] for i in rng.integers(0, 500, 50_000)]
    codes, dictionary = fsx.dictionary_encode(big)
    assert len(dictionary) == len(set(big)) and [dictionary[c] for c in codes[:1000]] == big[:1000]


def test_hll_estimator_matches_the_oracle_and_the_true_cardinality():
    for n in (1, 10, 1000, 40_000, 2_000_000):
        vals = (np.arange(n, dtype=np.int64) * 7919 + 12345)
        regs = orc.hll_registers(vals)
        est = fsx.hll_estimate(regs)
        assert abs(est - orc.hll_estimate(regs)) <= 1e-9 * est
        assert abs(est - n) <= max(1.0, 0.05 * n)                      # 3 sigma of a 2^12-register sketch
    a, b = orc.hll_registers(np.arange(0, 60_000, dtype=np.int32)), orc.hll_registers(np.arange(30_000, 90_000, dtype=np.int32))
    assert abs(fsx.hll_estimate(np.maximum(a, b)) - 90_000) <= 0.05 * 90_000   # sketches of shards merge by maximum
