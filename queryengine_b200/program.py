"""Tree -> flattened program serializer (the words qe_program_create takes).

Postfix walk of the expression tree; one 32-bit word per node,
``word = opcode | argument << 8`` (include/qegpu.h).  Opcodes 0..16 are the
ordinals of the reference's ``Function`` enum
(src/main/java/net/jhorstmann/queryengine/ast/Functions.kt:7-22).

The serializer does not rely on ``FunctionExpression.dataTypeNullable``: trees
built by hand in the reference's tests leave it unset
(src/test/java/net/jhorstmann/queryengine/ByteCodeCompilerTest.kt:28), so result
types are inferred by the library when the pipeline is bound to a table.
"""
from __future__ import annotations

import struct
from typing import List

import numpy as np

from .ast import (
    BooleanLiteralExpression, ColumnExpression, Expression, FunctionExpression,
    NumericLiteralExpression, StringLiteralExpression,
)

OP_LOAD_COL = 32
OP_CONST_F64 = 33
OP_CONST_BOOL = 34
OP_CONST_STR = 35
QE_CMP_IEEE = 1          # argument of CMP_LT / LE / GE / GT: IEEE ordering (include/qegpu.h)
_ORDERING = (11, 12, 13, 14)


class UnsupportedExpression(Exception):
    """The expression cannot run on the GPU path (e.g. a STRING literal without a dictionary)."""


def _emit(e: Expression, out: List[int], dictionary=None, ieee: bool = False) -> None:
    if isinstance(e, ColumnExpression):
        if e.index < 0 or e.index >= (1 << 24):
            raise ValueError("column index out of range")
        out.append(OP_LOAD_COL | (e.index << 8))
    elif isinstance(e, NumericLiteralExpression):
        bits = struct.unpack("<Q", struct.pack("<d", float(e.value)))[0]
        out += [OP_CONST_F64, bits & 0xFFFFFFFF, bits >> 32]
    elif isinstance(e, BooleanLiteralExpression):
        out.append(OP_CONST_BOOL | ((1 if e.value else 0) << 8))
    elif isinstance(e, StringLiteralExpression):
        # StringLiteralExpression (.../ast/Expressions.kt:29-33) as the code of the context's dictionary:
        # the same dictionary encodes the STRING columns, so equal codes <=> equal strings
        if dictionary is None:
            raise UnsupportedExpression("STRING literals need the context's StringDictionary")
        out += [OP_CONST_STR, dictionary.encode(e.value) & 0xFFFFFFFF]
    elif isinstance(e, FunctionExpression):
        for o in e.operands:
            _emit(o, out, dictionary, ieee)
        f = int(e.function)
        out.append(f | (QE_CMP_IEEE << 8) if ieee and f in _ORDERING else f)
    else:
        # IdentifierExpression must be resolved before evaluation
        # (src/main/java/net/jhorstmann/queryengine/evaluator/Interpreter.kt:9-11)
        raise TypeError(f"unexpected node {type(e).__name__} during evaluation")


def serialize(e: Expression, dictionary=None, ieee_order: bool = False) -> np.ndarray:
    """ieee_order: the ordering comparisons of Mode.CLOSURE_COMPILER
    (src/main/java/net/jhorstmann/queryengine/evaluator/ClosureCompiler.kt:127-130) instead of Double.compare's."""
    words: List[int] = []
    _emit(e, words, dictionary, ieee_order)
    return np.asarray(words, dtype=np.uint32)
