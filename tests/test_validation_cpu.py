"""Host logic of fsx/validation.py and the oracle restatements of the frequency / rank metrics (no GPU)."""
import math

import numpy as np
import pytest

from fsx import validation as V
from oracle import oracle as orc


def test_rule_catalogue_is_the_one_the_reference_lists(golden):
    cat = golden["G7"]["rule_catalogue"]
    assert {r["name"]: (r["predicate"], r["valueType"]) for r in cat} == V.RULE_CATALOGUE
    with pytest.raises(ValueError):
        V.Rule("HAS_TYPO")


def test_java_double_to_string_matches_the_frozen_value_strings(golden):
    g2 = golden["G2"]
    assert V.java_double_to_string(float(np.float32(140893.77))) == g2["expected"]["max_salary"]
    assert V.java_double_to_string(float(np.float32(52593.63))) == g2["expected"]["max_commission"]
    assert V.java_double_to_string(2020.0) == "2020.0" and V.java_double_to_string(0.0) == "0.0"
    # Double.toString switches to computerised scientific notation outside [1e-3, 1e7)
    assert [V.java_double_to_string(x) for x in (1e7, 9999999.0, 1e-3, 1.5e-5, -2.5, 123456789.125, float("nan"), float("-inf"))] == \
        ["1.0E7", "9999999.0", "0.001", "1.5E-5", "-2.5", "1.23456789125E8", "NaN", "-Infinity"]


def test_ingestion_policy_per_validation_type():
    assert V.ingestion_allowed("STRICT", "SUCCESS") and not V.ingestion_allowed("STRICT", "WARNING") and not V.ingestion_allowed("STRICT", "FAILURE")
    assert V.ingestion_allowed("WARNING", "WARNING") and not V.ingestion_allowed("WARNING", "FAILURE")
    assert V.ingestion_allowed("ALL", "FAILURE") and V.ingestion_allowed("NONE", "FAILURE")


def test_oracle_distinct_stats_against_first_principles():
    v = np.array([1.0, 1.0, 2.0, 3.0, 3.0, 3.0, 0.0, -0.0, np.nan, 7.5])
    valid = np.array([1, 1, 1, 1, 1, 1, 1, 1, 1, 0], bool)
    s = orc.distinct_stats(v, valid)
    assert (s["n_rows"], s["n_nonnull"], s["n_distinct"], s["n_unique"]) == (10, 9, 6, 4)      # 0.0 and -0.0 are two values
    p = np.array([2, 1, 3, 1, 1, 1]) / 9
    assert math.isclose(s["entropy"], -(p * np.log(p)).sum(), rel_tol=1e-15)
    assert orc.distinct_stats(np.zeros(0, np.float32))["n_distinct"] == 0


def test_oracle_quantiles_rank_definition():
    v = np.array([5, 1, 4, 2, 3], np.int32)
    assert list(orc.quantiles_exact(v, [0.0, 0.2, 0.21, 0.5, 0.99, 1.0])) == [1, 1, 2, 3, 5, 5]
    w = np.array([np.nan, 2.0, -1.0, np.inf])
    assert orc.quantiles_exact(w, [0.5])[0] == 2.0 and np.isnan(orc.quantiles_exact(w, [1.0])[0])    # NaN sorts above +inf
    assert np.isnan(orc.quantiles_exact(np.zeros(0), [0.5])[0])
