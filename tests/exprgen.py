"""Seeded random, well-typed expression trees over a TableSpec (test helper)."""
from __future__ import annotations

import numpy as np

from queryengine_b200.ast import (
    BooleanLiteralExpression, ColumnExpression, DataType, Expression, Function, FunctionExpression,
    NumericLiteralExpression, StringLiteralExpression,
)

STRINGS = ["a", "b", "", "lineitem", "N", "never in a column"]
LITERALS = [0.0, -0.0, 1.0, -1.0, 2.0, 0.5, 7.0, 100.0, -3.25, float("inf"), float("nan"), 1e-3, 24.0]


class ExprGen:
    def __init__(self, rng: np.random.Generator, fields):
        self.rng = rng
        self.dcols = [i for i, f in enumerate(fields) if f.type in (DataType.DOUBLE, DataType.INT64, DataType.DATE32)]
        self.bcols = [i for i, f in enumerate(fields) if f.type == DataType.BOOLEAN]
        self.scols = [i for i, f in enumerate(fields) if f.type == DataType.STRING]
        self.fields = fields

    def _col(self, i) -> ColumnExpression:
        return ColumnExpression(self.fields[i].name, i, self.fields[i].type)

    def double(self, depth: int) -> Expression:
        r = self.rng
        if depth <= 0 or r.random() < 0.2:
            if self.dcols and r.random() < 0.75:
                return self._col(int(r.choice(self.dcols)))
            return NumericLiteralExpression(float(LITERALS[int(r.integers(len(LITERALS)))]))
        pick = r.random()
        if pick < 0.70:
            f = [Function.ADD, Function.SUB, Function.MUL, Function.DIV, Function.MOD][int(r.integers(5))]
            return FunctionExpression(f, [self.double(depth - 1), self.double(depth - 1)])
        if pick < 0.85:
            f = Function.UNARY_MINUS if r.random() < 0.8 else Function.UNARY_PLUS
            return FunctionExpression(f, [self.double(depth - 1)])
        return FunctionExpression(Function.IF, [self.boolean(depth - 1), self.double(depth - 1), self.double(depth - 1)])

    def string(self, depth: int) -> Expression:
        """STRING-typed trees: columns, literals, IF (all the reference's TypeCheck allows, TypeCheck.kt:86-94)."""
        r = self.rng
        if depth <= 0 or r.random() < 0.4:
            if self.scols and r.random() < 0.7:
                return self._col(int(r.choice(self.scols)))
            return StringLiteralExpression(STRINGS[int(r.integers(len(STRINGS)))])
        return FunctionExpression(Function.IF, [self.boolean(depth - 1), self.string(depth - 1), self.string(depth - 1)])

    def boolean(self, depth: int) -> Expression:
        r = self.rng
        if self.scols and depth > 0 and r.random() < 0.2:      # string equality, TypeCheck.kt:65-71
            f = Function.CMP_EQ if r.random() < 0.5 else Function.CMP_NE
            return FunctionExpression(f, [self.string(depth - 1), self.string(depth - 1)])
        if depth <= 0 or r.random() < 0.1:
            if self.bcols and r.random() < 0.8:
                return self._col(int(r.choice(self.bcols)))
            if depth <= 0 and not self.bcols:
                return FunctionExpression(Function.CMP_LT, [self.double(0), self.double(0)])
            return BooleanLiteralExpression(bool(r.random() < 0.5))
        pick = r.random()
        if pick < 0.45:
            f = [Function.CMP_LT, Function.CMP_LE, Function.CMP_GE, Function.CMP_GT, Function.CMP_EQ,
                 Function.CMP_NE][int(r.integers(6))]
            return FunctionExpression(f, [self.double(depth - 1), self.double(depth - 1)])
        if pick < 0.80:
            f = Function.AND if r.random() < 0.5 else Function.OR
            return FunctionExpression(f, [self.boolean(depth - 1), self.boolean(depth - 1)])
        if pick < 0.90:
            return FunctionExpression(Function.NOT, [self.boolean(depth - 1)])
        if pick < 0.95:
            f = Function.CMP_EQ if r.random() < 0.5 else Function.CMP_NE
            return FunctionExpression(f, [self.boolean(depth - 1), self.boolean(depth - 1)])
        return FunctionExpression(Function.IF, [self.boolean(depth - 1), self.boolean(depth - 1), self.boolean(depth - 1)])
