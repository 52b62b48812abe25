"""Feature validation on top of the GPU statistics: the host-side mirror of hsfs' Rule / Expectation /
FeatureGroupValidation objects (reference: notebooks/featurestore/hsfs/data_validation/
feature_validation_python.ipynb -- rule catalogue :218-243, expectations :304-315, validation results :498,
validation types :741-750).  hsfs evaluates the rules with Deequ analyzers over a Spark DataFrame; here every
metric a rule asserts on comes from the CUDA kernels behind the C ABI (fsx_stats_describe, fsx_stats_distinct,
fsx_stats_quantiles, fsx_stats_pearson, fsx_stats_histogram).  Nothing is computed on the CPU except the
comparison of one number per (feature, rule) against its bounds.

Rules on strings (HAS_DATATYPE, HAS_PATTERN, IS_CONTAINED_IN) and on column pairs by row (IS_LESS_THAN ...) are outside
the numeric hot path and raise NotImplementedError.
"""
from __future__ import annotations

import math
import time
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence

from . import api

# name -> (predicate, valueType) exactly as connection.get_rules() lists them (notebook :218-243)
RULE_CATALOGUE = {
    "HAS_SIZE": ("VALUE", "Integral"), "HAS_MEAN": ("VALUE", "Fractional"), "HAS_DATATYPE": ("ACCEPTED_TYPE", "String"),
    "HAS_SUM": ("VALUE", "Fractional"), "IS_LESS_THAN": ("LEGAL_VALUES", "Fractional"),
    "IS_GREATER_THAN_OR_EQUAL_TO": ("LEGAL_VALUES", "Fractional"), "HAS_STANDARD_DEVIATION": ("VALUE", "Fractional"),
    "HAS_PATTERN": ("PATTERN", "String"), "HAS_NUMBER_OF_DISTINCT_VALUES": ("VALUE", "Integral"), "HAS_MAX": ("VALUE", "Fractional"),
    "IS_CONTAINED_IN": ("LEGAL_VALUES", "String"), "IS_NON_NEGATIVE": ("VALUE", "Fractional"), "IS_POSITIVE": ("VALUE", "Boolean"),
    "HAS_MUTUAL_INFORMATION": ("LEGAL_VALUES", "Fractional"), "HAS_UNIQUE_VALUE_RATIO": ("VALUE", "Fractional"),
    "IS_GREATER_THAN": ("LEGAL_VALUES", "Fractional"), "HAS_COMPLETENESS": ("VALUE", "Fractional"), "HAS_ENTROPY": ("VALUE", "Fractional"),
    "HAS_MIN": ("VALUE", "Fractional"), "HAS_UNIQUENESS": ("VALUE", "Fractional"), "HAS_DISTINCTNESS": ("VALUE", "Fractional"),
    "HAS_CORRELATION": ("LEGAL_VALUES", "Fractional"), "HAS_APPROX_QUANTILE": ("VALUE", "Fractional"),
    "HAS_APPROX_COUNT_DISTINCT": ("VALUE", "Fractional"), "IS_LESS_THAN_OR_EQUAL_TO": ("LEGAL_VALUES", "Fractional"),
}
_UNSUPPORTED = {"HAS_DATATYPE", "HAS_PATTERN", "IS_CONTAINED_IN", "IS_LESS_THAN", "IS_LESS_THAN_OR_EQUAL_TO", "IS_GREATER_THAN",
                "IS_GREATER_THAN_OR_EQUAL_TO"}


def java_double_to_string(x: float) -> str:
    """java.lang.Double.toString: the `value` strings of a validation result ('140893.765625', '2020.0', '1.0E10')"""
    if math.isnan(x):
        return "NaN"
    if math.isinf(x):
        return "Infinity" if x > 0 else "-Infinity"
    if x == 0:
        return "-0.0" if math.copysign(1.0, x) < 0 else "0.0"
    from decimal import Decimal
    t = Decimal(repr(abs(x))).as_tuple()                   # shortest digits that round-trip, like Java's algorithm
    mant = "".join(map(str, t.digits)).rstrip("0") or "0"
    e10 = len(t.digits) + t.exponent - 1                   # decimal exponent of the first digit
    lead = len("".join(map(str, t.digits))) - len("".join(map(str, t.digits)).lstrip("0"))
    mant, e10 = (mant.lstrip("0") or "0"), e10 - lead
    sign = "-" if x < 0 else ""
    if 1e-3 <= abs(x) < 1e7:
        if e10 >= 0:
            ip, fp = mant[:e10 + 1].ljust(e10 + 1, "0"), mant[e10 + 1:]
        else:
            ip, fp = "0", "0" * (-e10 - 1) + mant
        return f"{sign}{ip}.{fp or '0'}"
    return f"{sign}{mant[0]}.{mant[1:] or '0'}E{e10}"


@dataclass
class Rule:
    name: str
    level: str = "WARNING"                                 # WARNING | ERROR
    min: Optional[float] = None
    max: Optional[float] = None
    pattern: Optional[str] = None
    accepted_type: Optional[str] = None
    legal_values: Optional[Sequence] = None                # HAS_CORRELATION: [other feature]
    quantile: float = 0.5                                  # HAS_APPROX_QUANTILE

    def __post_init__(self):
        if self.name not in RULE_CATALOGUE:
            raise ValueError(f"unknown rule {self.name}")
        if self.level not in ("WARNING", "ERROR"):
            raise ValueError(f"rule level {self.level}")

    def to_dict(self):
        d = {"name": self.name, "level": self.level}
        if self.min is not None:
            d["min"] = float(self.min)
        if self.max is not None:
            d["max"] = float(self.max)
        return d


@dataclass
class Expectation:
    name: str
    features: List[str]
    rules: List[Rule]
    description: str = ""

    def to_dict(self):
        return {"name": self.name, "description": self.description, "features": list(self.features), "rules": [r.to_dict() for r in self.rules]}


@dataclass
class ValidationResult:
    status: str                                            # SUCCESS | FAILURE
    message: str
    value: str
    feature: str
    rule: Rule

    def to_dict(self):
        return {"feature": self.feature, "message": self.message, "rule": self.rule.to_dict(), "status": self.status, "value": self.value}


@dataclass
class ExpectationResult:
    expectation: Expectation
    results: List[ValidationResult]

    @property
    def status(self):
        # the frozen rejected-upsert output reports Failure for an unmet WARNING-level rule (notebook :741): an
        # expectation is SUCCESS only if every rule held; an unmet rule is WARNING only when asked to grade by level
        return "SUCCESS" if all(r.status == "SUCCESS" for r in self.results) else "FAILURE"

    def to_dict(self):
        return {"expectation": self.expectation.to_dict(), "results": [r.to_dict() for r in self.results], "status": self.status}


@dataclass
class FeatureGroupValidation:
    expectation_results: List[ExpectationResult]
    validation_time: int = field(default_factory=lambda: int(time.time() * 1000))

    @property
    def status(self):
        return "SUCCESS" if all(e.status == "SUCCESS" for e in self.expectation_results) else "FAILURE"

    def to_dict(self):
        return {"validationTime": self.validation_time, "expectationResults": [e.to_dict() for e in self.expectation_results], "status": self.status}


def ingestion_allowed(validation_type: str, status: str) -> bool:
    """STRICT: only SUCCESS; WARNING: SUCCESS or WARNING; ALL: always; NONE: validation is not performed (notebook :741-750)."""
    vt = validation_type.upper()
    if vt in ("ALL", "NONE"):
        return True
    if vt == "STRICT":
        return status == "SUCCESS"
    if vt == "WARNING":
        return status in ("SUCCESS", "WARNING")
    raise ValueError(f"validation type {validation_type}")


class _Metrics:
    """per-feature metrics, each computed on the GPU at most once"""

    def __init__(self, ctx: api.Context, columns: Dict[str, api.Column]):
        self.ctx, self.columns = ctx, columns
        self._desc, self._dist, self._quant, self._corr = {}, {}, {}, {}

    def describe(self, f):
        if f not in self._desc:
            names = [n for n in self.columns if n not in self._desc]
            for n, m in zip(names, self.ctx.describe([self.columns[n] for n in names])):   # one pass for every column still missing
                self._desc[n] = m
        return self._desc[f]

    def distinct(self, f):
        if f not in self._dist:
            self._dist[f] = self.ctx.distinct_stats(self.columns[f])
        return self._dist[f]

    def quantile(self, f, q):
        if (f, q) not in self._quant:
            self._quant[(f, q)] = float(self.ctx.quantiles(self.columns[f], [q])[0][0])
        return self._quant[(f, q)]

    def correlation(self, a, b):
        if (a, b) not in self._corr:
            self._corr[(a, b)] = float(self.ctx.pearson([self.columns[a], self.columns[b]])[0, 1])
        return self._corr[(a, b)]

    def mutual_information(self, a, b):
        """I(A;B) = H(A) + H(B) - H(A,B) in nats (Deequ MutualInformation): three entropies from fsx_stats_distinct, the joint
        one over the combined key of the two columns (fsx_column_combine_keys: null-free integer columns)."""
        if (a, b) not in self._corr:
            ca, cb = self.columns[a], self.columns[b]
            if ca.info()[2] or cb.info()[2] or ca.info()[0] not in (api.I32, api.I64) or cb.info()[0] not in (api.I32, api.I64):
                raise NotImplementedError("HAS_MUTUAL_INFORMATION: null-free integer (or dictionary-encoded) columns only")
            joint = self.ctx.combine_keys(ca, cb)
            self._corr[(a, b)] = self.distinct(a).entropy + self.distinct(b).entropy - self.ctx.distinct_stats(joint).entropy
        return self._corr[(a, b)]

    def fraction_at_least(self, f, lo):
        """fraction of the rows whose value lies in [lo, max] or is null (Deequ's compliance of `COALESCE(col, ok) >= lo`):
        one histogram bucket with explicit bounds"""
        m = self.describe(f)
        if m.n_rows == 0:
            return 1.0
        nulls = m.n_rows - m.n_nonnull
        if m.n_nonnull == 0 or not (m.max >= lo):
            return nulls / m.n_rows
        counts, _, _ = self.ctx.histogram([self.columns[f]], 1, [lo], [m.max])
        return (float(counts[0][0]) + nulls) / m.n_rows

    def value(self, feature: str, rule: Rule) -> float:
        n = rule.name
        if n in _UNSUPPORTED:
            raise NotImplementedError(f"{n}: outside the numeric hot path (strings, row-wise column comparisons)")
        if n in ("HAS_MIN", "HAS_MAX", "HAS_MEAN", "HAS_SUM", "HAS_STANDARD_DEVIATION", "HAS_SIZE", "HAS_COMPLETENESS"):
            m = self.describe(feature)
            nn = m.n_nonnull
            return {"HAS_MIN": m.min, "HAS_MAX": m.max, "HAS_MEAN": m.mean, "HAS_SUM": m.sum, "HAS_SIZE": float(m.n_rows),
                    "HAS_STANDARD_DEVIATION": math.sqrt(m.m2 / nn) if nn else float("nan"),      # Deequ StandardDeviation: population
                    "HAS_COMPLETENESS": nn / m.n_rows if m.n_rows else 1.0}[n]
        if n in ("HAS_NUMBER_OF_DISTINCT_VALUES", "HAS_APPROX_COUNT_DISTINCT", "HAS_DISTINCTNESS", "HAS_UNIQUENESS", "HAS_UNIQUE_VALUE_RATIO", "HAS_ENTROPY"):
            d = self.distinct(feature)
            nn = d.n_nonnull
            return {"HAS_NUMBER_OF_DISTINCT_VALUES": float(d.n_distinct), "HAS_APPROX_COUNT_DISTINCT": float(d.n_distinct),
                    "HAS_DISTINCTNESS": d.n_distinct / nn if nn else 0.0, "HAS_UNIQUENESS": d.n_unique / nn if nn else 0.0,
                    "HAS_UNIQUE_VALUE_RATIO": d.n_unique / d.n_distinct if d.n_distinct else 0.0, "HAS_ENTROPY": d.entropy}[n]
        if n == "HAS_APPROX_QUANTILE":
            return self.quantile(feature, rule.quantile)
        if n == "HAS_CORRELATION":
            if not rule.legal_values:
                raise ValueError("HAS_CORRELATION needs legal_values=[other feature]")
            return self.correlation(feature, rule.legal_values[0])
        if n == "HAS_MUTUAL_INFORMATION":
            if not rule.legal_values:
                raise ValueError("HAS_MUTUAL_INFORMATION needs legal_values=[other feature]")
            return self.mutual_information(feature, rule.legal_values[0])
        if n == "IS_NON_NEGATIVE":
            return self.fraction_at_least(feature, 0.0)
        if n == "IS_POSITIVE":
            return self.fraction_at_least(feature, 5e-324)
        raise NotImplementedError(n)


def validate(ctx: api.Context, columns: Dict[str, api.Column], expectations: Sequence[Expectation],
             validation_time: Optional[int] = None) -> FeatureGroupValidation:
    """economy_fg.validate(df): every (feature, rule) of every expectation, metrics from the GPU."""
    metrics = _Metrics(ctx, columns)
    out = []
    for ex in expectations:
        results = []
        for rule in ex.rules:
            for feature in ex.features:
                if feature not in columns:
                    raise KeyError(f"expectation {ex.name}: feature {feature} is not in the dataframe")
                v = metrics.value(feature, rule)
                lo = rule.min if rule.min is not None else (1.0 if rule.name in ("IS_NON_NEGATIVE", "IS_POSITIVE") and rule.max is None else None)
                ok = not math.isnan(v) and (lo is None or v >= lo) and (rule.max is None or v <= rule.max)
                sv = java_double_to_string(v)
                results.append(ValidationResult("SUCCESS" if ok else "FAILURE",
                                                "Success" if ok else f"Value: {sv} does not meet the constraint requirement! {rule.name}",
                                                sv, feature, rule))
        out.append(ExpectationResult(ex, results))
    return FeatureGroupValidation(out) if validation_time is None else FeatureGroupValidation(out, validation_time)
