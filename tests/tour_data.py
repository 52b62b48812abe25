"""Synthetic tour-shaped tables (the real football CSVs are not in the
reference checkout -- featurestore_tour/.gitignore:25 ignores data/*).  Schemas follow the Raw* case
classes (ComputeFeatures.scala:39-89); seed 1."""
import os

import numpy as np


def write_tour_csvs(out_dir: str, seed: int = 1, null_team_rows: int = 0):
    rng = np.random.default_rng(seed)
    os.makedirs(out_dir, exist_ok=True)
    n_teams = 50
    team_id = np.arange(1, n_teams + 1)
    with open(os.path.join(out_dir, "teams.csv"), "w") as f:
        f.write("team_budget,team_id,team_name,team_owner,team_position\n")
        pos = rng.permutation(team_id)
        for t, p in zip(team_id, pos):
            f.write(f"{rng.uniform(1e6, 1e8)!r},{t},\"Team, {t}\",owner_{t},{p}\n")
    with open(os.path.join(out_dir, "players.csv"), "w") as f:
        f.write("age,rating,team_id,worth\n")
        rows = [(rng.integers(18, 39), rng.lognormal(4, 0.3), t, rng.lognormal(13, 1.0)) for t in team_id for _ in range(25)]
        for i in rng.permutation(len(rows)):
            a, r, t, w = rows[i]
            f.write(f"{a},{float(r)!r},{t},{float(w)!r}\n")
        for _ in range(null_team_rows):          # rows with an empty team_id: the null-key group (dropped by the join)
            f.write(f"{rng.integers(18, 39)},{float(rng.lognormal(4, 0.3))!r},,{float(rng.lognormal(13, 1.0))!r}\n")
    with open(os.path.join(out_dir, "attendances.csv"), "w") as f:
        f.write("attendance,team_id,year\n")
        for t in team_id:
            for y in range(2010, 2020):
                f.write(f"{float(rng.uniform(5e3, 9e4)) + 0.5!r},{t},{y}\n")
    with open(os.path.join(out_dir, "season_scores.csv"), "w") as f:
        f.write("position,team_id,year\n")
        for y in range(2010, 2020):
            for t, p in zip(team_id, rng.permutation(team_id)):
                f.write(f"{p},{t},{y}\n")
    with open(os.path.join(out_dir, "games.csv"), "w") as f:
        f.write("away_team_id,home_team_id,score\n")
        for h in team_id:
            for a in team_id:
                if a != h:
                    f.write(f"{a},{h},{rng.integers(0, 6)}\n")
    return out_dir
