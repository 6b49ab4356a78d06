"""File-backed tables on the GPU: the same data written as Arrow IPC, Parquet and the
reference's CSV dialect, scanned / filtered / aggregated through FileTable, against the oracle fed with the
same Python values.  Needs a B200."""
import numpy as np
import pyarrow as pa
import pyarrow.parquet as pq
import pytest

from oracle_binding import OraclePlan, OracleTable
from queryengine_b200.ast import AggregationFunction as A, DataType, Function, StringLiteralExpression, agg, col, fn
from queryengine_b200.engine import Context, Field, GpuGroupByOperator, GpuPipelineOperator
from queryengine_b200.source import FileTable

pytestmark = pytest.mark.gpu
D, B, S = DataType.DOUBLE, DataType.BOOLEAN, DataType.STRING


@pytest.fixture(scope="module")
def ctx():
    c = Context(0)
    yield c
    c.close()


def _rows(n, seed=5):
    rng = np.random.default_rng(seed)
    flags = ["A", "N", "R"]
    rows = []
    for i in range(n):
        rows.append([None if rng.random() < 0.1 else float(np.round(rng.uniform(-50, 50), 2)),
                     None if rng.random() < 0.05 else flags[int(rng.integers(3))],
                     None if rng.random() < 0.1 else bool(rng.random() < 0.5),
                     float(int(rng.integers(0, 100)))])
    return rows


FIELDS = [Field("x", D, True), Field("flag", S, True), Field("ok", B, True), Field("q", D, True)]


def _write(tmp_path, rows, kind):
    cols = list(zip(*rows)) if rows else [[], [], [], []]
    t = pa.table([pa.array(cols[0], pa.float64()), pa.array(cols[1], pa.string()), pa.array(cols[2], pa.bool_()),
                  pa.array(cols[3], pa.float64())], names=[f.name for f in FIELDS])
    p = str(tmp_path / ("t." + kind))
    if kind == "arrow":
        with pa.OSFile(p, "wb") as f, pa.ipc.new_file(f, t.schema) as w:
            for b in t.to_batches(max_chunksize=1000):         # several record batches, not multiples of 8 rows
                w.write_batch(b)
    elif kind == "parquet":
        pq.write_table(t, p, row_group_size=777, use_dictionary=True)
    else:
        with open(p, "w") as f:
            f.write(",".join(fl.name for fl in FIELDS) + "\n")
            for r in rows:
                f.write(",".join("" if v is None else ("true" if v is True else "false" if v is False else repr(v) if isinstance(v, float) else v)
                                 for v in r) + "\n")
    return p


@pytest.mark.parametrize("kind", ["arrow", "parquet", "csv"])
def test_file_table_scan_filter_groupby_match_oracle(ctx, tmp_path, kind):
    rows = _rows(5003)
    path = _write(tmp_path, rows, kind)
    ft = FileTable(ctx, path, FIELDS if kind == "csv" else None)
    assert [(f.name, f.type) for f in ft.fields] == [(f.name, f.type) for f in FIELDS]
    # Table.getScanOperator(projection): the named columns, in that order
    got = ft.getScanOperator(["flag", "x"]).map()
    assert got == [[r[1], r[0]] for r in rows]
    otable = OracleTable.from_rows([f.type for f in FIELDS], rows)
    x, flag, ok, q = (col(i, f.type) for i, f in enumerate(FIELDS))
    t = ft.load([f.name for f in FIELDS])
    filt = fn(Function.AND, fn(Function.CMP_NE, flag, StringLiteralExpression("N")), fn(Function.OR, ok, fn(Function.CMP_LT, x, 0.0)))
    projs = [fn(Function.MUL, x, q), flag, ok]
    got = GpuPipelineOperator(t, filt, projs, None).map()
    want = OraclePlan(otable, [0, 1, 2, 3], filter=filt, projections=projs).rows()
    assert got == want
    got = GpuGroupByOperator(t, None, [flag, agg(A.COUNT, x), agg(A.MAX, q), agg(A.MIN, x)]).map()
    want = OraclePlan(otable, [0, 1, 2, 3], projections=[flag, x, q, x], aggregates=[A.COUNT, A.MAX, A.MIN], group_count=1).rows()
    assert got == want
    with pytest.raises(ValueError):
        ft.getScanOperator(["nope"])
    ft.close()


def test_empty_file(ctx, tmp_path):
    path = _write(tmp_path, [], "arrow")
    ft = FileTable(ctx, path)
    assert ft.getScanOperator(["x", "q"]).map() == []
    ft.close()
