"""The tour job restated on the CPU oracle: ComputeFeatures.scala:101-329 step by step with Spark
2.4 semantics (groupBy.sum / count / inner join / typed map in float32 / hsfs primary-key joins),
plus the statistics the feature groups are saved with.  Test infrastructure only."""
import numpy as np
import pandas as pd

from oracle import oracle as orc


def _read(path):
    """CSV -> {column: (values, valid mask or None)} with Spark 2.4 inference restated independently:
    Integer -> Long -> Double -> String over the non-empty cells."""
    import csv
    import re
    with open(path, newline="") as f:
        rows = list(csv.reader(f))
    header, body = rows[0], [r for r in rows[1:] if r]
    out = {}
    for ci, name in enumerate(header):
        cells = [r[ci] for r in body]
        valid = np.array([c != "" for c in cells], bool)
        present = [c for c in cells if c != ""]
        if present and all(re.fullmatch(r"[+-]?\d+", c) for c in present):
            vals = np.array([int(c) if c != "" else 0 for c in cells], dtype=np.int64)
            if np.abs(vals).max() < 2**31:
                vals = vals.astype(np.int32)
        else:
            try:
                vals = np.array([float(c) if c != "" else 0.0 for c in cells], dtype=np.float64)
            except ValueError:
                vals = np.array(cells, dtype=object)
        out[name] = (vals, None if valid.all() else valid)
    return out


def _aggregate_by_team(raw, specs):
    keys, kvalid = raw["team_id"]
    gid, uniq, null_gid = orc.group_ids(keys, kvalid)
    G = len(uniq)
    cnt = orc.group_count(gid, G)
    keep = np.array([g for g in np.argsort(uniq, kind="stable") if g != null_gid], dtype=np.int64)   # inner join drops the null group
    keep = keep[np.argsort(uniq[keep], kind="stable")]
    out = {"team_id": uniq[keep]}
    for src, avg_name, sum_name in specs:
        vals, vvalid = raw[src]
        a = orc.group_agg(vals, gid, G, vvalid)
        s = a["sum_i64"] if vals.dtype.kind == "i" else a["sum_f64"]
        avg, sm = orc.finalize_avg(s[keep], cnt[keep])
        out[avg_name], out[sum_name] = avg, sm
    return out


def run(input_dir):
    """-> {feature group name: {column: ndarray}} in case-class field order."""
    fgs = {}
    games = _read(f"{input_dir}/games.csv")
    fgs["games_features"] = {c: games[c][0] for c in ("away_team_id", "home_team_id", "score")}
    fgs["games_features_hudi_tour"] = fgs["games_features"]
    s = _aggregate_by_team(_read(f"{input_dir}/season_scores.csv"), [("position", "average_position", "sum_position")])
    fgs["season_scores_features"] = {k: s[k] for k in ("team_id", "average_position", "sum_position")}
    a = _aggregate_by_team(_read(f"{input_dir}/attendances.csv"), [("attendance", "average_attendance", "sum_attendance")])
    fgs["attendances_features"] = {k: a[k] for k in ("team_id", "average_attendance", "sum_attendance")}
    p = _aggregate_by_team(_read(f"{input_dir}/players.csv"),
                           [("age", "average_player_age", "sum_player_age"), ("rating", "average_player_rating", "sum_player_rating"),
                            ("worth", "average_player_worth", "sum_player_worth")])
    fgs["players_features"] = {k: p[k] for k in ("team_id", "average_player_rating", "average_player_age", "average_player_worth",
                                                 "sum_player_rating", "sum_player_age", "sum_player_worth")}
    teams = _read(f"{input_dir}/teams.csv")
    fgs["teams_features"] = {"team_budget": orc.d2f(teams["team_budget"][0]), "team_id": teams["team_id"][0],
                             "team_position": teams["team_position"][0]}
    pl, tm, at = fgs["players_features"], fgs["teams_features"], fgs["attendances_features"]
    l1, r1 = orc.inner_join(pl["team_id"], tm["team_id"])
    k1 = pl["team_id"][l1]
    l2, r2 = orc.inner_join(k1, at["team_id"])
    fgs["tour_training_dataset_test"] = {"average_player_age": pl["average_player_age"][l1][l2], "team_budget": tm["team_budget"][r1][l2],
                                         "average_attendance": at["average_attendance"][r2]}
    return fgs


def statistics(cols: dict):
    """describe + 20-bin histogram + pearson of a feature group, column order preserved."""
    names = list(cols)
    out = {"columns": {}}
    for n in names:
        m = orc.moments(cols[n])
        rc, counts, edges = orc.histogram(cols[n], 20)
        ds = orc.distinct_stats(cols[n])
        out["columns"][n] = dict(count=m.n_rows, mean=m.mean, minimum=m.min, maximum=m.max, sum=m.sum,
                                 stdDev=float(np.sqrt(m.m2 / m.n_nonnull)), distinct=ds["n_distinct"], uniqueness=ds["n_unique"] / ds["n_nonnull"],
                                 entropy=ds["entropy"], percentiles=orc.quantiles_exact(cols[n], np.arange(1, 101) / 100.0).tolist(),
                                 histogram=counts[:rc].tolist(), edges=edges[:rc + 1].tolist() if rc > 1 else [edges[0], edges[0]])
    out["corr"] = orc.pearson([cols[n] for n in names]) if len(names) >= 2 and len(cols[names[0]]) > 1 else None
    return out
