"""The BASELINE.json configurations as expression trees and schemas, in one place:
bench.py runs them, queryengine_b200/aot.py compiles their kernels into the library ahead of time, the full-size
tests check them.  Pure host bookkeeping over queryengine_b200.ast; nothing here touches a device.
"""
from __future__ import annotations

from typing import List, Tuple

from .ast import AggregationFunction as A
from .ast import DataType, Expression, Function, col, fn

SEED = 42
SELECTIVITIES = (0.01, 0.50, 0.99)

D, S, I64 = DataType.DOUBLE, DataType.STRING, DataType.INT64

# (name, type, nullable) per column
Schema = List[Tuple[str, DataType, bool]]


def single_column_schema(dtype: DataType = I64) -> Schema:
    return [("col", dtype, False)]


def single_column_filter(s: float, dtype: DataType = I64) -> Expression:
    """configs[1]: `col < k` with k = floor(s * 2^40) over values uniform in [0, 2^40)."""
    return fn(Function.CMP_LT, col(0, dtype, "col"), float(int(s * 2 ** 40)))


def q6_schema() -> Schema:
    return [("l_quantity", D, False), ("l_extendedprice", D, False), ("l_discount", D, False), ("l_shipdate", D, False)]


def q6_filter() -> Expression:
    """configs[2]: the TPC-H Q6 predicate over (quantity, extendedprice, discount, shipdate), all f64."""
    return fn(Function.AND, fn(Function.AND, fn(Function.AND, fn(Function.AND,
              fn(Function.CMP_GE, col(3), 8766.0), fn(Function.CMP_LT, col(3), 9131.0)),
              fn(Function.CMP_GE, col(2), 0.05)), fn(Function.CMP_LE, col(2), 0.07)), fn(Function.CMP_LT, col(0), 24.0))


def q6_projection() -> Expression:
    return fn(Function.MUL, col(1), col(2))


def tree_schema() -> Schema:
    return [(n, D, True) for n in "abcd"]


def tree_expr() -> Expression:
    """configs[3]: the depth-8 nullable arithmetic tree (fixed here, evaluated by the oracle in the tests)."""
    a, b, c_, d = col(0), col(1), col(2), col(3)
    inner = fn(Function.DIV, fn(Function.MUL, fn(Function.ADD, a, b), fn(Function.SUB, c_, d)),
               fn(Function.ADD, fn(Function.MUL, a, 0.5), fn(Function.SUB, fn(Function.MOD, c_, 7.0), fn(Function.UNARY_MINUS, b))))
    return fn(Function.IF, fn(Function.CMP_GT, inner, 0.0),
              fn(Function.ADD, fn(Function.MUL, inner, fn(Function.SUB, a, d)), fn(Function.DIV, c_, fn(Function.ADD, b, 1.5))),
              fn(Function.SUB, fn(Function.UNARY_MINUS, inner), fn(Function.MUL, d, 0.25)))


def lineitem_schema() -> Schema:
    return [("l_quantity", D, False), ("l_extendedprice", D, False), ("l_discount", D, False), ("l_tax", D, False),
            ("l_shipdate", D, False)]


def lineitem_filter() -> Expression:
    """configs[4]: multi-column AND / OR filter over (qty, extendedprice, discount, tax, shipdate)."""
    qty, ep, disc, tax, sd = (col(i) for i in range(5))
    return fn(Function.OR,
              fn(Function.AND, fn(Function.AND, fn(Function.CMP_GE, sd, 8766.0), fn(Function.CMP_LT, sd, 9131.0)),
                 fn(Function.CMP_LT, qty, 24.0)),
              fn(Function.AND, fn(Function.CMP_GE, disc, 0.05), fn(Function.CMP_LE, tax, 0.02)))


def lineitem_projections() -> List[Expression]:
    qty, ep, disc, tax, sd = (col(i) for i in range(5))
    net = fn(Function.MUL, ep, fn(Function.SUB, 1.0, disc))
    gross = fn(Function.MUL, fn(Function.MUL, ep, fn(Function.SUB, 1.0, disc)), fn(Function.ADD, 1.0, tax))
    return [net, gross]


def q1_schema() -> Schema:
    return [("l_returnflag", S, False), ("l_linestatus", S, False), ("l_quantity", D, False), ("l_extendedprice", D, False),
            ("l_discount", D, False), ("l_tax", D, False), ("l_shipdate", D, False)]


def q1_groupby():
    """TPC-H Q1 shape: (filter, keys, aggregate inputs, functions)."""
    rf, ls, qty, ep, disc, tax, sd = col(0, S), col(1, S), col(2), col(3), col(4), col(5), col(6)
    net = fn(Function.MUL, ep, fn(Function.SUB, 1.0, disc))
    gross = fn(Function.MUL, net, fn(Function.ADD, 1.0, tax))
    return (fn(Function.CMP_LE, sd, 10471.0), [rf, ls], [qty, ep, net, gross, qty, ep, disc, rf],
            [A.SUM, A.SUM, A.SUM, A.SUM, A.AVG, A.AVG, A.AVG, A.COUNT])
