"""Expression tree, host-side mirror of the reference's AST for the hot path.

Same names, members and ordinals as the reference so that tests read like the
reference's own (src/test/java/net/jhorstmann/queryengine/evaluator/CompilerTest.kt):

* ``DataType``            - src/main/java/net/jhorstmann/queryengine/data/Schema.kt:3-5
                            (+ INT64 / DATE32 extensions that have no reference counterpart,
                            SURVEY.md section 8a row T)
* ``Function``            - .../ast/Functions.kt:7-22 (ordinals are the opcode numbers of the
                            flattened program, include/qegpu.h)
* ``AggregationFunction`` - .../ast/Functions.kt:24-26
* expression node classes - .../ast/Expressions.kt:6-62
* ``Mode``                - .../evaluator/Compiler.kt:5-7, plus the new ``GPU`` member

This module is pure host bookkeeping; it evaluates nothing.
"""
from __future__ import annotations

import enum
from dataclasses import dataclass, field
from typing import List, Optional, Sequence, Tuple


class DataType(enum.IntEnum):
    STRING = 0
    DOUBLE = 1
    BOOLEAN = 2
    # extensions: physical column types that widen exactly to DOUBLE on load
    INT64 = 3
    DATE32 = 4


class FunctionType(enum.IntEnum):
    LOGIC = 0
    ARITHMETIC = 1
    COMPARISON = 2


class Function(enum.IntEnum):
    AND = 0
    OR = 1
    IF = 2
    NOT = 3
    UNARY_MINUS = 4
    UNARY_PLUS = 5
    MUL = 6
    DIV = 7
    MOD = 8
    ADD = 9
    SUB = 10
    CMP_LT = 11
    CMP_LE = 12
    CMP_GE = 13
    CMP_GT = 14
    CMP_EQ = 15
    CMP_NE = 16

    @property
    def arity(self) -> int:
        return _ARITY[self]

    @property
    def type(self) -> FunctionType:
        return _FTYPE[self]


_ARITY = {
    Function.AND: 2, Function.OR: 2, Function.IF: 3, Function.NOT: 1,
    Function.UNARY_MINUS: 1, Function.UNARY_PLUS: 1,
    Function.MUL: 2, Function.DIV: 2, Function.MOD: 2, Function.ADD: 2, Function.SUB: 2,
    Function.CMP_LT: 2, Function.CMP_LE: 2, Function.CMP_GE: 2, Function.CMP_GT: 2,
    Function.CMP_EQ: 2, Function.CMP_NE: 2,
}
_FTYPE = {f: (FunctionType.LOGIC if f <= Function.NOT else
              FunctionType.ARITHMETIC if f <= Function.SUB else FunctionType.COMPARISON)
          for f in Function}


class AggregationFunction(enum.IntEnum):
    MIN = 0
    MAX = 1
    SUM = 2
    COUNT = 3
    AVG = 4
    ANY = 5
    ALL = 6


class Mode(enum.IntEnum):
    INTERPRETER = 0
    CLOSURE_COMPILER = 1
    BYTECODE_COMPILER = 2
    GPU = 3


class Expression:
    """Base class (sealed class Expression, Expressions.kt:6-9)."""
    __slots__ = ()


@dataclass(frozen=True)
class ColumnExpression(Expression):
    name: str
    index: int
    dataType: DataType


@dataclass(frozen=True)
class NumericLiteralExpression(Expression):
    value: float

    @property
    def dataType(self) -> DataType:
        return DataType.DOUBLE


@dataclass(frozen=True)
class BooleanLiteralExpression(Expression):
    value: bool

    @property
    def dataType(self) -> DataType:
        return DataType.BOOLEAN


@dataclass(frozen=True)
class StringLiteralExpression(Expression):
    value: str

    @property
    def dataType(self) -> DataType:
        return DataType.STRING


@dataclass(frozen=True)
class FunctionExpression(Expression):
    function: Function
    operands: Tuple[Expression, ...]
    dataTypeNullable: Optional[DataType] = None

    def __init__(self, function: Function, operands: Sequence[Expression],
                 dataTypeNullable: Optional[DataType] = None):
        object.__setattr__(self, "function", Function(function))
        object.__setattr__(self, "operands", tuple(operands))
        object.__setattr__(self, "dataTypeNullable", dataTypeNullable)
        if len(self.operands) != self.function.arity:
            raise ValueError(f"{self.function.name} takes {self.function.arity} operands, got {len(self.operands)}")


@dataclass(frozen=True)
class AggregationFunctionExpression(Expression):
    """AggregationFunctionExpression(function, operands) (Expressions.kt:47-57); one operand."""
    function: AggregationFunction
    operands: Tuple[Expression, ...]

    def __init__(self, function: AggregationFunction, operands: Sequence[Expression]):
        object.__setattr__(self, "function", AggregationFunction(function))
        object.__setattr__(self, "operands", tuple(operands))
        if len(self.operands) != 1:
            raise ValueError("aggregation functions take one operand")


# ---- small construction helpers used by tests and the bench harness -------

def col(index: int, dtype: DataType = DataType.DOUBLE, name: Optional[str] = None) -> ColumnExpression:
    return ColumnExpression(name or f"c{index}", index, dtype)


def lit(v) -> Expression:
    if isinstance(v, bool):
        return BooleanLiteralExpression(v)
    if isinstance(v, str):
        return StringLiteralExpression(v)
    return NumericLiteralExpression(float(v))


def fn(function: Function, *ops) -> FunctionExpression:
    return FunctionExpression(function, [o if isinstance(o, Expression) else lit(o) for o in ops])


def agg(function: AggregationFunction, operand) -> AggregationFunctionExpression:
    return AggregationFunctionExpression(function, [operand if isinstance(operand, Expression) else lit(operand)])


def walk(e: Expression):
    """Post-order traversal."""
    if isinstance(e, FunctionExpression):
        for o in e.operands:
            yield from walk(o)
    yield e


def referenced_columns(e: Expression) -> List[int]:
    seen: List[int] = []
    for n in walk(e):
        if isinstance(n, ColumnExpression) and n.index not in seen:
            seen.append(n.index)
    return seen
