import sys, time, os
sys.path.insert(0, os.getcwd())
from queryengine_b200 import capi
from queryengine_b200.ast import AggregationFunction as A, DataType, col
from queryengine_b200.engine import ColumnarTable, Context, Field, GroupByPipeline
ctx = Context(0)
D = DataType.DOUBLE
n = 600_000_000
t = ColumnarTable(ctx, [Field("k", D), Field("x", D)], n)
t.generate(1, capi.QE_GEN_UNIFORM_REAL, 5, 0.0, 1.0)
for card in (4, 1024, 2000, 100000):
    t.generate(0, capi.QE_GEN_UNIFORM_INT, 6, 0, card - 1)
    pipe = GroupByPipeline(t, None, [col(0)], [col(1), col(1)], [A.SUM, A.COUNT], max_groups=max(65536, 2 * card))
    res = pipe.run(); pipe.run(res)
    m0, n0 = res.kernel_stats()
    t0 = time.perf_counter()
    for _ in range(3): pipe.run(res)
    wall = (time.perf_counter() - t0) / 3 * 1e3
    res.num_rows
    m1, n1 = res.kernel_stats()
    print(card, "variant", capi.lib().qe_result_group_variant(res.handle), "kernel_ms", round((m1 - m0) / max(1, n1 - n0), 3), "wall_ms per call", round(wall, 3))
    res.close()
