"""BASELINE.json config 1: the featurestore_tour job end to end through the host-side mirror
(libfsx_tour.so -> libfsx.so) on synthetic tour-shaped tables, against the same job restated on the
CPU oracle: every feature group's rows and the statistics it is saved with."""
import json
import os

import numpy as np
import pandas as pd
import pytest

import tour_oracle
from fsx import tour
from helpers import assert_close
from tour_data import write_tour_csvs

pytestmark = pytest.mark.gpu

FLOAT32_FROM_DOUBLE_SUM = {"average_attendance", "sum_attendance", "average_player_rating", "average_player_worth",
                           "sum_player_rating", "sum_player_worth"}


def _load(out_dir, name):
    df = pd.read_csv(os.path.join(out_dir, f"{name}_1.csv"))
    stats = json.load(open(os.path.join(out_dir, f"{name}_1.statistics.json")))
    return df, stats


@pytest.fixture(scope="module", params=[0, 7], ids=["clean", "null_team_ids"])
def job(request, tmp_path_factory):
    d = tmp_path_factory.mktemp("tour")
    inp = write_tour_csvs(str(d / "in"), seed=1, null_team_rows=request.param)
    out = str(d / "out")
    tour.run(inp, out)
    return inp, out, tour_oracle.run(inp)


def test_every_feature_group_is_written_with_the_reference_schema(job):
    _, out, ref = job
    for name, cols in ref.items():
        df, _ = _load(out, name)
        assert list(df.columns) == list(cols), name                # case-class field order
        assert len(df) == len(next(iter(cols.values()))), name     # row counts / join cardinalities: exact
    assert os.path.exists(os.path.join(out, "season_features_on_demand_1.statistics.json"))
    meta = json.load(open(os.path.join(out, "season_scores_features_1.meta.json")))
    assert meta["onlineEnabled"] is True and meta["primaryKeys"] == ["team_id"]
    meta = json.load(open(os.path.join(out, "games_features_hudi_tour_1.meta.json")))
    assert meta["timeTravelFormat"] == "HUDI" and meta["partitionKeys"] == ["score"]


def test_feature_values_match_the_oracle(job):
    _, out, ref = job
    for name, cols in ref.items():
        df, _ = _load(out, name)
        if name == "tour_training_dataset_test":
            df = df.sort_values(list(df.columns)).reset_index(drop=True)
            cols = pd.DataFrame(cols).sort_values(list(cols)).reset_index(drop=True)
        for c in df.columns:
            got, want = df[c].to_numpy(), np.asarray(cols[c])
            if want.dtype.kind == "i":
                assert np.array_equal(got, want), f"{name}.{c}"
            elif c in FLOAT32_FROM_DOUBLE_SUM:
                # float32 cast of an order-dependent FP64 sum: equal or the neighbouring float
                ulp = np.abs(got.astype(np.float32).view(np.int32).astype(np.int64) - want.astype(np.float32).view(np.int32))
                assert ulp.max() <= 1 and (ulp == 0).mean() > 0.9, f"{name}.{c}"
            else:
                assert np.array_equal(got.astype(np.float32), want.astype(np.float32)), f"{name}.{c} must be bit-exact"


def test_statistics_match_the_oracle(job):
    _, out, ref = job
    for name in ("games_features", "season_scores_features", "attendances_features", "players_features", "teams_features",
                 "tour_training_dataset_test"):
        df, stats = _load(out, name)
        cols = {c: df[c].to_numpy().astype(np.asarray(ref[name][c]).dtype) for c in df.columns}    # dtypes of the case classes
        want = tour_oracle.statistics(cols)
        assert [c["column"] for c in stats["columns"]] == list(cols)
        for i, c in enumerate(stats["columns"]):
            w = want["columns"][c["column"]]
            assert c["count"] == w["count"] and c["numDistinctValuesExact"] == w["distinct"] and abs(c["approximateNumDistinctValues"] - w["distinct"]) <= max(1, 0.05 * w["distinct"])
            assert c["minimum"] == w["minimum"] and c["maximum"] == w["maximum"]
            assert c["dataType"] == ("Integral" if cols[c["column"]].dtype.kind == "i" else "Fractional")
            for k in ("mean", "sum", "stdDev", "uniqueness", "entropy"):
                assert_close(c[k], w[k], what=f"{name}.{c['column']}.{k}")
            assert c["approxPercentiles"] == w["percentiles"], f"{name}.{c['column']} percentiles are order statistics: exact"
            assert [h["count"] for h in c["histogram"]] == w["histogram"], f"{name}.{c['column']} histogram bins"
            assert sum(h["count"] for h in c["histogram"]) == w["count"]
            got_corr = np.array([np.nan if x["correlation"] is None else x["correlation"] for x in c["correlations"]])
            assert_close(got_corr, want["corr"][i], what=f"{name}.{c['column']} correlations")


def test_type_mismatch_fails_like_getas(tmp_path):
    d = write_tour_csvs(str(tmp_path / "in"))
    p = os.path.join(d, "attendances.csv")
    txt = open(p).read().replace(".", "")                 # integer-looking attendance: Spark infers int, getAs[Double] throws
    open(p, "w").write(txt)
    with pytest.raises(RuntimeError, match="attendance"):
        tour.run(d, str(tmp_path / "out"))
