"""Host logic of the tour mirror without a GPU: CSV reading with Spark-2.4 schema inference
(ComputeFeatures.scala:104 `option("inferSchema","true")`) and Main.parseInput."""
import os

import pytest

from fsx import tour
from tour_data import write_tour_csvs


def test_tour_library_loads_and_binds():
    assert os.path.exists(tour.TOUR_LIB_PATH)
    tour.load()


def test_schema_inference_on_tour_tables(tmp_path):
    d = write_tour_csvs(str(tmp_path))
    cols, n = tour.infer_schema(os.path.join(d, "players.csv"))
    assert cols == [("age", "int", False), ("rating", "double", False), ("team_id", "int", False), ("worth", "double", False)] and n == 1250
    cols, n = tour.infer_schema(os.path.join(d, "teams.csv"))
    assert [c[:2] for c in cols] == [("team_budget", "double"), ("team_id", "int"), ("team_name", "string"), ("team_owner", "string"),
                                     ("team_position", "int")] and n == 50
    assert tour.infer_schema(os.path.join(d, "games.csv"))[1] == 2450
    assert [c[1] for c in tour.infer_schema(os.path.join(d, "attendances.csv"))[0]] == ["double", "int", "int"]
    assert [c[1] for c in tour.infer_schema(os.path.join(d, "season_scores.csv"))[0]] == ["int", "int", "int"]


def test_schema_inference_rules(tmp_path):
    p = tmp_path / "t.csv"
    p.write_text("a,b,c,d,e,f,g\n"
                 "1,1,1,x,,2147483647,\"q,uoted\"\n"
                 "2,3000000000,1.5,2,,2147483648,1\n"
                 ",4,-2e3,3,,5,2\n")
    cols, n = tour.infer_schema(str(p))
    # int; int->bigint; int->double; string sticks; all-empty -> string; overflowing int -> bigint; quoted text -> string
    assert cols == [("a", "int", True), ("b", "bigint", False), ("c", "double", False), ("d", "string", False),
                    ("e", "string", True), ("f", "bigint", False), ("g", "string", False)] and n == 3
    # a literal NA forces the column to string (no nullValue option is set in the reference)
    p.write_text("x,y\n1.5,NA\n2.5,3\n")
    assert [c[1] for c in tour.infer_schema(str(p))[0]] == ["double", "string"]
    p.write_text("x,y\n1,2\n3\n")
    with pytest.raises(RuntimeError):
        tour.infer_schema(str(p))                         # ragged row
    with pytest.raises(RuntimeError):
        tour.infer_schema(str(tmp_path / "missing.csv"))


def test_parse_input_mirrors_main():
    assert tour.parse_input("hdfs://namenode/Projects/x/data", "demo") == "hdfs://namenode/Projects/x/data"
    assert tour.parse_input("TestJob/data", "demo") == "hdfs:///Projects/demo/TestJob/data"    # Main.scala:64
    assert tour.parse_input("/tmp/data", "demo") == "/tmp/data"
