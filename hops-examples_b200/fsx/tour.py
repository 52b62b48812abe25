"""ctypes binding of libfsx_tour.so: the host-side mirror of the reference job
(hops-examples_b200/host/tour.hpp; ComputeFeatures.scala + Main.scala) driven the way a JVM would
drive it through JNI.  All compute happens in libfsx kernels; nothing here falls back to the CPU."""
from __future__ import annotations

import ctypes as C
import os

from . import _lib

TOUR_LIB_PATH = os.path.join(os.path.dirname(_lib.LIB_PATH), "libfsx_tour.so")
_tour = None


def load() -> C.CDLL:
    global _tour
    if _tour is None:
        _lib.load()  # libfsx.so first (RTLD_GLOBAL) so the tour library binds to the same instance
        if not os.path.exists(TOUR_LIB_PATH):
            raise RuntimeError(f"libfsx_tour.so not found at {TOUR_LIB_PATH}: run `make -C hops-examples_b200`")
        t = C.CDLL(TOUR_LIB_PATH)
        t.fsx_tour_run.restype = C.c_int32
        t.fsx_tour_run.argtypes = [C.c_char_p, C.c_char_p, C.c_int32, C.c_int32, C.c_char_p, C.c_size_t]
        t.fsx_tour_infer_schema.restype = C.c_int32
        t.fsx_tour_infer_schema.argtypes = [C.c_char_p, C.c_char_p, C.c_size_t, C.POINTER(C.c_int64)]
        t.fsx_tour_parse_input.restype = C.c_int32
        t.fsx_tour_parse_input.argtypes = [C.c_char_p, C.c_char_p, C.c_char_p, C.c_size_t]
        t.fsx_tour_read_csv_column.restype = C.c_int32
        t.fsx_tour_read_csv_column.argtypes = [C.c_char_p, C.c_int32, C.c_int64, C.c_void_p, C.c_void_p, C.POINTER(C.c_int32), C.POINTER(C.c_int64)]
        t.fsx_tour_profile_csv.restype = C.c_int32
        t.fsx_tour_profile_csv.argtypes = [C.c_char_p, C.c_int32, C.c_char_p, C.c_size_t, C.POINTER(C.c_int64)]
        t.fsx_tour_delta_files.restype = C.c_int32
        t.fsx_tour_delta_files.argtypes = [C.c_char_p, C.c_char_p, C.c_size_t]
        t.fsx_tour_profile_parquet.restype = C.c_int32
        t.fsx_tour_profile_parquet.argtypes = [C.c_char_p, C.c_int32, C.c_char_p, C.c_size_t, C.POINTER(C.c_int64), C.c_char_p, C.c_size_t]
        _tour = t
    return _tour


def run(input_dir: str, output_dir: str, device: int = 0, quiet: bool = True) -> None:
    """Main.main(--input input_dir): the seven ComputeFeatures calls in the reference's order; every
    saved feature group becomes <output_dir>/<name>_<version>.csv + .statistics.json."""
    err = C.create_string_buffer(2048)
    rc = load().fsx_tour_run(input_dir.encode(), output_dir.encode(), device, int(quiet), err, len(err))
    if rc != 0:
        raise RuntimeError(f"fsx_tour_run failed ({rc}): {err.value.decode(errors='replace')}")


def infer_schema(csv_path: str):
    """-> ([(name, type, has_nulls)...], nrows) with Spark 2.4 CSV inference rules (CPU, no GPU needed)."""
    out = C.create_string_buffer(1 << 16)
    n = C.c_int64()
    rc = load().fsx_tour_infer_schema(csv_path.encode(), out, len(out), C.byref(n))
    if rc != 0:
        raise RuntimeError(out.value.decode(errors="replace"))
    cols = []
    for item in out.value.decode().split(","):
        name, typ = item.rsplit(":", 1)
        cols.append((name, typ.rstrip("?"), typ.endswith("?")))
    return cols, n.value


def parse_input(path: str, project: str) -> str:
    out = C.create_string_buffer(4096)
    load().fsx_tour_parse_input(path.encode(), project.encode(), out, len(out))
    return out.value.decode()


def profile_csv(csv_path: str, device: int = 0) -> dict:
    """Statistics of every column of a headered CSV as the Deequ-style JSON of the job (ComputeFeatures::profileCsv):
    numeric columns through fsx_stats_profile, string columns dictionary-coded on the host and profiled as categorical."""
    import json
    size = 1 << 20
    while True:
        out = C.create_string_buffer(size)
        need = C.c_int64()
        rc = load().fsx_tour_profile_csv(csv_path.encode(), device, out, len(out), C.byref(need))
        if rc != 0:
            raise RuntimeError(out.value.decode(errors="replace"))
        if need.value <= size:
            return json.loads(out.value.decode())
        size = need.value + 16


def delta_files(table_dir: str):
    """the part files a Delta table's transaction log leaves active, in the order they were added (host code)"""
    out = C.create_string_buffer(1 << 20)
    if load().fsx_tour_delta_files(table_dir.encode(), out, len(out)) != 0:
        raise RuntimeError(out.value.decode(errors="replace"))
    return [p for p in out.value.decode().split("\n") if p]


def profile_parquet(path: str, device: int = 0):
    """Statistics of the columns of one .parquet file -- or of a Delta table directory, all its active part files as one
    table -- decoded on the GPU (ComputeFeatures::profileParquet)
    -> (statistics dict, names of the columns libfsx declined to decode)"""
    import json
    size = 1 << 20
    while True:
        out, skipped = C.create_string_buffer(size), C.create_string_buffer(1 << 16)
        need = C.c_int64()
        rc = load().fsx_tour_profile_parquet(path.encode(), device, out, len(out), C.byref(need), skipped, len(skipped))
        if rc != 0:
            raise RuntimeError(out.value.decode(errors="replace"))
        if need.value <= size:
            return json.loads(out.value.decode()), [s for s in skipped.value.decode().split(",") if s]
        size = need.value + 16


def read_csv_column(csv_path: str, col: int, cap: int = 1 << 22):
    """one column as the HOST reader (host/csv.hpp) parses it -> (type name, cells as uint64 bit patterns, valid bool array)"""
    import numpy as np
    cells, valid = np.zeros(cap, np.uint64), np.zeros(cap, np.uint8)
    ty, n = C.c_int32(), C.c_int64()
    rc = load().fsx_tour_read_csv_column(csv_path.encode(), col, cap, cells.ctypes.data, valid.ctypes.data, C.byref(ty), C.byref(n))
    if rc != 0:
        raise RuntimeError(f"fsx_tour_read_csv_column failed ({rc})")
    return {1: "int", 2: "bigint", 3: "double", 4: "string"}[ty.value], cells[:n.value], valid[:n.value].astype(bool)
