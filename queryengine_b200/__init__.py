"""queryengine_b200 -- B200-native filter/scan hot path of jhorstmann/queryengine.

Only what the path needs: the CUDA kernels + C ABI (csrc/, built into
libqegpu.so) and the host-side mirror of the reference's operator/expression
interface.  There is no CPU fallback: without the built library every compute
entry point raises.
"""
from .ast import (  # noqa: F401
    AggregationFunction, BooleanLiteralExpression, ColumnExpression, DataType, Expression,
    Function, FunctionExpression, FunctionType, Mode, NumericLiteralExpression,
    StringLiteralExpression, col, fn, lit,
)

__version__ = "0.1.0"
