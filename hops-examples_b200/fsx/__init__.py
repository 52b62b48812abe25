"""fsx -- Python binding of libfsx (include/fsx.h): the B200-native engine behind the
featurestore_tour feature-engineering job (groupBy/agg, equi-joins, feature-group statistics).

Import path: the package lives in ``hops-examples_b200/fsx`` (the directory name of the project
is not a valid Python identifier), so callers put ``hops-examples_b200`` on ``sys.path`` --
``__graft_entry__`` and ``tests/conftest.py`` do this.
"""
from .api import (  # noqa: F401
    AVG, COUNT, COUNT_STAR, F32, F64, GEN_IOTA, GEN_MIX, GEN_NORMALISH, GEN_UNIFORM_INT, I32, I64,
    FULL_OUTER, INNER, LEFT_ANTI, LEFT_OUTER, LEFT_SEMI, RIGHT_OUTER, MAX, MIN, SUM, Column, Context, Describe, FsxError, Join, Table,
    dtype_code, pack_validity, unpack_validity, dictionary_encode, hll_estimate, parquet_schema,
    ROUTE_AUTO, ROUTE_DENSE, ROUTE_PARTIAL, ROUTE_ROWS, SHUFFLE, BROADCAST,
)
from ._lib import LIB_PATH, SIGNATURES, load  # noqa: F401
