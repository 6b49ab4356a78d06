"""CPU-side checks of the drop-in boundary (no GPU): the C-ABI library loads and
exports every symbol include/qegpu.h declares; programs are flattened, bound and
type-checked like the reference's TypeCheck; generated pipeline code compiles for
sm_100a with NVRTC; without a GPU the compute entry points fail loudly."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from queryengine_b200 import capi
from queryengine_b200.ast import (
    AggregationFunction, BooleanLiteralExpression, ColumnExpression, DataType, Function, FunctionExpression,
    NumericLiteralExpression, StringLiteralExpression, col, fn, lit,
)
from queryengine_b200.build import build
from queryengine_b200.program import UnsupportedExpression, serialize

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
D, B, I, T = DataType.DOUBLE, DataType.BOOLEAN, DataType.INT64, DataType.DATE32


@pytest.fixture(scope="module", autouse=True)
def _built():
    build()


def test_library_exports_every_declared_symbol():
    header = open(os.path.join(ROOT, "include", "qegpu.h")).read()
    declared = set(re.findall(r"QE_API\s+[\w\s\*]+?\b(qe_\w+)\s*\(", header))
    assert len(declared) >= 30
    L = C.CDLL(capi.LIB_PATH)
    for name in declared:
        assert hasattr(L, name), f"{name} declared in qegpu.h but not exported"
    assert declared == set(capi.SYMBOLS), declared ^ set(capi.SYMBOLS)
    assert capi.lib().qe_version().decode().count(".") == 2


def test_ffm_binding_source_names_real_entry_points_with_the_right_arity():
    """bindings/ffm/QeGpuFfm.java (Panama FFM, JDK 22+; source only, no JDK here): every downcall handle names a function
    the library exports, with as many arguments as include/qegpu.h declares."""
    src = open(os.path.join(ROOT, "bindings", "ffm", "QeGpuFfm.java")).read()
    header = re.sub(r"/\*.*?\*/", "", open(os.path.join(ROOT, "include", "qegpu.h")).read(), flags=re.S)
    handles = re.findall(r'h\("(qe_\w+)",\s*FunctionDescriptor\.of\(([^;]*?)\)\);', src, flags=re.S)
    assert len(handles) >= 14
    L = C.CDLL(capi.LIB_PATH)
    for name, layouts in handles:
        assert hasattr(L, name), name
        decl = re.search(r"QE_API\s+[\w\s\*]+?\b" + name + r"\s*\(([^;]*?)\)\s*;", header, flags=re.S)
        assert decl, name
        params = [x for x in decl.group(1).split(",") if x.strip() and x.strip() != "void"]
        nargs = len([x for x in layouts.split(",") if x.strip()]) - 1          # the first layout is the return value
        assert nargs == len(params), (name, nargs, params)


def test_serializer_postfix_words():
    e = fn(Function.CMP_LT, col(3), 5.0)
    w = serialize(e)
    assert w[0] == (32 | (3 << 8)) and w[1] == 33 and int(w[4]) == int(Function.CMP_LT)
    assert np.array([w[2], w[3]], dtype=np.uint32).view(np.float64)[0] == 5.0
    assert list(serialize(lit(True))) == [34 | (1 << 8)]
    with pytest.raises(UnsupportedExpression):
        serialize(StringLiteralExpression("t"))          # no dictionary given
    with pytest.raises(ValueError):
        FunctionExpression(Function.ADD, [lit(1.0)])


def _generate(schema, filter_, projs, aggs=(), steps=0, vec=0, compile_=True, staging=-1, dictionary=None):
    L = capi.lib()
    descs = (capi.ColDesc * len(schema))(*[capi.ColDesc(int(t), 1 if n else 0) for t, n in schema])
    exprs = ([filter_] if filter_ is not None else []) + list(projs)
    words = [serialize(e, dictionary) for e in exprs]
    ptrs = (C.c_void_p * len(words))(*[w.ctypes.data for w in words])
    lens = (C.c_size_t * len(words))(*[w.size for w in words])
    aggarr = (C.c_int32 * max(1, len(aggs)))(*[int(a) for a in aggs])
    src = C.create_string_buffer(1 << 16)
    nbytes = C.c_size_t(0)
    rc = L.qe_debug_generate(descs, len(schema), ptrs, lens, 1 if filter_ is not None else 0, len(projs),
                             aggarr, len(aggs), steps, vec, staging, 1 if compile_ else 0, src, len(src), None, 0, C.byref(nbytes))
    return rc, src.value.decode(), nbytes.value, L.qe_last_error().decode()


Q6_SCHEMA = [(D, False)] * 4      # quantity, extendedprice, discount, shipdate
Q6_FILTER = fn(Function.AND, fn(Function.AND, fn(Function.AND, fn(Function.AND,
               fn(Function.CMP_GE, col(3), 8766.0), fn(Function.CMP_LT, col(3), 9131.0)),
               fn(Function.CMP_GE, col(2), 0.05)), fn(Function.CMP_LE, col(2), 0.07)),
               fn(Function.CMP_LT, col(0), 24.0))
Q6_PROJ = fn(Function.MUL, col(1), col(2))


def test_generated_code_compiles_for_sm100a():
    rc, src, nbytes, err = _generate([(I, False)], fn(Function.CMP_LT, col(0, I), 1e6), [col(0, I)])
    assert rc == 0, err
    # the literal is a kernel parameter: the code names its slot, not its value
    assert "qe_filter_ring_body<P>" in src and "((qe_i64)in.c0 < p.lit_hi[0])" in src and nbytes > 1000
    assert "TILE_ROWS = 4480" in src and "NCW = 20" in src and "row.c0 = qe_stage1_i64(stage + 0, wrow0, c);" in src
    rc, src, nbytes, err = _generate([(I, False)], fn(Function.CMP_LT, col(0, I), 1e6), [col(0, I)], staging=0)
    assert rc == 0, err
    assert "qe_filter_pipelined_body<P>" in src and nbytes > 1000
    rc, src, nbytes, err = _generate([(I, False)], fn(Function.CMP_LT, col(0, I), 1e6), [col(0, I)], staging=1)
    assert rc == 0, err
    assert "qe_bulk_g2s(stage + 0, (const char *)p.col_values[0] + row0 * 8" in src and "qe_filter_staged_body<P>" in src
    rc, src, nbytes, err = _generate(Q6_SCHEMA, Q6_FILTER, [Q6_PROJ], [AggregationFunction.SUM])
    assert rc == 0, err
    assert "qe_agg_body<P>" in src and "qe_acc_add<2>" in src
    # nullable arithmetic tree, dense projection with validity
    tree = fn(Function.DIV, fn(Function.MUL, fn(Function.ADD, col(0), col(1)), fn(Function.SUB, col(2), col(3))),
              fn(Function.ADD, fn(Function.MUL, col(0), 0.5),
                 fn(Function.SUB, fn(Function.MOD, col(2), 7.0), fn(Function.UNARY_MINUS, col(1)))))
    rc, src, nbytes, err = _generate([(D, False), (D, True), (D, True), (D, True)], None, [tree])
    assert rc == 0, err
    assert "qe_project_body<P>" in src and "NBITCOLS = 1" in src and "qe_mod_k(" in src
    rc, src, nbytes, err = _generate([(D, False), (D, True), (D, True), (D, True)], None, [tree], staging=2)
    assert rc == 0, err
    assert "qe_project_ring_body<P>" in src and "qe_project_ring_kernel" in src
    # every tile shape the runtime may pick
    for steps in (1, 2, 4, 8):
        for vec in (2, 4):
            for staging in (0, 1, 2):
                rc, src, nbytes, err = _generate([(D, True), (B, True), (T, False)],
                                                 fn(Function.OR, col(1, B), fn(Function.CMP_GE, col(2, T), 9000.0)),
                                                 [fn(Function.IF, col(1, B), col(0), 1.0), col(1, B), col(2, T)],
                                                 steps=steps, vec=vec, staging=staging)
                assert rc == 0, (steps, vec, staging, err)
                assert ("qe_filter_staged_body<P>" in src) == (staging == 1)
                assert ("qe_filter_ring_body<P>" in src) == (staging == 2)


def test_literals_are_kernel_parameters():
    """`col < 10` and `col < 11` are ONE kernel (the cache key held the literal bits): the generated source
    and therefore the cubin cache key depend on the literal's class only -- NaN / zero / integral / MOD-able."""
    def src_of(schema, filt, projs, **kw):
        rc, src, _, err = _generate(schema, filt, projs, compile_=False, **kw)
        assert rc == 0, err
        return src
    a = src_of([(I, False)], fn(Function.CMP_LT, col(0, I), 10.0), [col(0, I)])
    b = src_of([(I, False)], fn(Function.CMP_LT, col(0, I), 11.0), [col(0, I)])
    assert a == b and "10" not in a.split("struct P")[1].split("eval(")[1].split("return")[0]
    # another class -> another kernel: a non-integral literal against an integer column folds EQ to `false`
    c = src_of([(I, False)], fn(Function.CMP_EQ, col(0, I), 10.0), [col(0, I)])
    d = src_of([(I, False)], fn(Function.CMP_EQ, col(0, I), 10.5), [col(0, I)])
    assert c != d and "== p.lit_lo[0]" in c and "= false;" in d
    # f64 columns: a plain literal takes the one-instruction comparison, NaN and zeros the total-order one
    e = src_of([(D, False)], fn(Function.CMP_LT, col(0), 3.0), [col(0)])
    f = src_of([(D, False)], fn(Function.CMP_LT, col(0), 4.5), [col(0)])
    g = src_of([(D, False)], fn(Function.CMP_LT, col(0), 0.0), [col(0)])
    h = src_of([(D, False)], fn(Function.CMP_LT, col(0), float("nan")), [col(0)])
    assert e == f and "qe_cmp_lt_k(in.c0, p.lit_f64[0])" in e
    assert "qe_cmp_lt(in.c0, p.lit_f64[0])" in g and "qe_cmp_lt(in.c0, p.lit_f64[0])" in h
    # equal literals share a slot, so a repeated subtree is still evaluated once (common subexpressions)
    sub = fn(Function.MOD, fn(Function.MUL, col(0), 0.5), 7.0)
    t = src_of([(D, False)], None, [fn(Function.ADD, sub, sub)])
    assert t.count("qe_mod_k(") == 1 and "p.lit_abs[1]" in t and "p.lit_inv[1]" in t
    # more distinct literals than parameter slots: the rest stay in the code
    e2 = col(0)
    for i in range(30):
        e2 = fn(Function.ADD, e2, float(i) + 0.25)
    u = src_of([(D, False)], None, [e2])
    assert "p.lit_f64[23]" in u and "p.lit_f64[24]" not in u and "qe_bits_f64((qe_i64)0x" in u
    # string literals: the dictionary code is a parameter as well
    from queryengine_b200.engine import StringDictionary
    dic = StringDictionary()
    S = DataType.STRING
    s1 = src_of([(S, True)], fn(Function.CMP_EQ, col(0, S), lit("a")), [col(0, S)], dictionary=dic)
    s2 = src_of([(S, True)], fn(Function.CMP_EQ, col(0, S), lit("another")), [col(0, S)], dictionary=dic)
    assert s1 == s2 and "(qe_i32)p.lit_lo[0]" in s1


def test_validity_the_filter_decides_is_written_as_runs():
    """Ring filter: an output that cannot be null in a kept row -- it is null only if a column is that the predicate needs
    non-null (FilterOperator.kt:20 keeps non-null true only; Kleene AND is true only if both sides are,
    Interpreter.kt:54-72) -- has no per-row validity bit in the generated code (NBITCOLS), its bitmap is P::implied_col."""
    def src_of(schema, filt, projs):
        rc, src, _, err = _generate(schema, filt, projs, compile_=True)
        assert rc == 0, err
        return src
    two = [(D, True), (D, True)]
    a = src_of(two, fn(Function.CMP_LT, col(0), 5.0), [col(0), fn(Function.MUL, col(0), 2.0)])
    assert "NBITCOLS = 0" in a and "NIMPLIED = 2" in a and "case 1: return p.out_valid[1];" in a
    # the other column is not read by the predicate: its validity stays a bit column
    b = src_of(two, fn(Function.CMP_LT, col(0), 5.0), [col(0), col(1)])
    assert "NBITCOLS = 1" in b and "NIMPLIED = 1" in b and "case 0: return p.out_valid[1];" in b.split("bitcol(")[1]
    # AND passes the promise on, OR and IF do not; NOT and arithmetic do
    c = src_of(two, fn(Function.AND, fn(Function.CMP_LT, col(0), 5.0), fn(Function.NOT, fn(Function.CMP_GT, fn(Function.ADD, col(1), 1.0), 2.0))),
               [fn(Function.ADD, col(0), col(1))])
    assert "NBITCOLS = 0" in c and "NIMPLIED = 1" in c
    d = src_of(two, fn(Function.OR, fn(Function.CMP_LT, col(0), 5.0), fn(Function.CMP_GT, col(1), 2.0)), [col(0), col(1)])
    assert "NBITCOLS = 2" in d and "NIMPLIED = 0" in d
    e = src_of(two, fn(Function.CMP_LT, col(0), 5.0), [fn(Function.IF, fn(Function.CMP_GT, col(0), 1.0), col(0), 0.0)])
    assert "NBITCOLS = 1" in e and "NIMPLIED = 0" in e
    # the other filter bodies write every bit themselves
    rc, f, _, err = _generate(two, fn(Function.CMP_LT, col(0), 5.0), [col(0)], staging=0)
    assert rc == 0 and "NBITCOLS = 1" in f and "NIMPLIED" not in f, err


@pytest.mark.parametrize("seed", range(40))
def test_implied_validity_is_sound_against_the_oracle(seed):
    """Soundness of the static analysis behind the runs of ones (implied_valid_outputs, qe_codegen.cpp): for random
    well-typed filter / projection trees over nullable columns, every output the generated code treats as "never null in
    a kept row" must be non-null in every row the ORACLE's FilterOperator keeps (no GPU needed: the generated source
    names the outputs, the oracle evaluates the rows)."""
    import sys
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from exprgen import ExprGen
    from oracle_binding import OraclePlan
    from parity import random_table
    from queryengine_b200.ast import Mode
    rng = np.random.default_rng(9000 + seed)
    spec = random_table(rng, 600, ["DDDB", "DDBB", "DdDB", "DDDD"][seed % 4], special_rate=0.2)
    g = ExprGen(rng, spec.fields)
    # filters that are conjunctions of comparisons more often than not: those are the ones that promise something
    parts = [g.boolean(2) for _ in range(int(rng.integers(1, 4)))]
    filt = parts[0]
    for q in parts[1:]:
        filt = fn(Function.AND, filt, q) if rng.random() < 0.8 else fn(Function.OR, filt, q)
    projs = [g.double(int(rng.integers(0, 4))) for _ in range(3)] + [g.boolean(2)]
    schema = [(f.type, f.nullable) for f in spec.fields]
    rc, src, _, err = _generate(schema, filt, projs, compile_=False)
    assert rc == 0, err
    body = src.split("implied_col(")[1].split("}")[0]
    implied = [int(k) for k in re.findall(r"return p\.out_valid\[(\d+)\];", body)]
    n_ref, vals, valid = OraclePlan(spec.oracle(), list(range(len(spec.fields))), filter=filt, projections=projs,
                                    mode=Mode.BYTECODE_COMPILER).columns()
    for k in implied:
        assert valid[k][:n_ref].all(), (k, seed)
    # and the analysis is not vacuous over the forty seeds: a plain conjunction over the columns an output reads qualifies
    if seed == 0:
        rc, src2, _, err = _generate(schema, fn(Function.AND, fn(Function.CMP_LT, col(0), 1.0), fn(Function.CMP_GT, col(1), -1.0)),
                                     [fn(Function.ADD, col(0), col(1))], compile_=False)
        assert rc == 0 and "NIMPLIED = 1" in src2, err


def test_closure_compiler_mode_is_a_comparison_argument():
    """Mode.CLOSURE_COMPILER (IEEE ordering, M/evaluator/ClosureCompiler.kt:127-130) travels as QE_CMP_IEEE in the
    argument of CMP_LT / LE / GE / GT: another kernel than the total-order one, the same kernel for an integer column
    against an ordinary literal (no NaN, no -0.0 on either side), and nothing else takes an argument."""
    from queryengine_b200.program import QE_CMP_IEEE

    def gen(schema, filt, ieee):
        words = [serialize(filt, None, ieee_order=ieee), serialize(col(0, schema[0][0]))]
        L = capi.lib()
        descs = (capi.ColDesc * len(schema))(*[capi.ColDesc(int(t), 1 if n else 0) for t, n in schema])
        ptrs = (C.c_void_p * 2)(*[w.ctypes.data for w in words])
        lens = (C.c_size_t * 2)(*[w.size for w in words])
        src = C.create_string_buffer(1 << 16)
        nbytes = C.c_size_t(0)
        rc = L.qe_debug_generate(descs, len(schema), ptrs, lens, 1, 1, (C.c_int32 * 1)(), 0, 0, 0, -1, 1, src, len(src),
                                 None, 0, C.byref(nbytes))
        assert rc == 0 and nbytes.value > 1000, L.qe_last_error().decode()
        return src.value.decode()
    lt = fn(Function.CMP_LT, col(0), col(1))
    a, b = gen([(D, True), (D, True)], lt, False), gen([(D, True), (D, True)], lt, True)
    assert "qe_cmp_lt(in.c0, in.c1)" in a and "(in.c0 < in.c1)" in b and "qe_cmp_lt(" not in b.split("struct P")[1]
    ge = fn(Function.CMP_GE, col(0), 3.0)                       # NaN >= 3.0: true in the total order, false in IEEE
    a, b = gen([(D, False)], ge, False), gen([(D, False)], ge, True)
    assert "qe_cmp_ge_k(in.c0, p.lit_f64[0])" in a and "(in.c0 >= p.lit_f64[0])" in b
    ik = fn(Function.CMP_LT, col(0, I), 10.0)
    assert "((qe_i64)in.c0 < p.lit_hi[0])" in gen([(I, False)], ik, True)
    # `a < b` in both orderings inside one tree: two values, not one common subexpression
    words = np.concatenate([serialize(lt, None, ieee_order=False), serialize(lt, None, ieee_order=True),
                            np.array([int(Function.CMP_NE)], np.uint32)])
    descs = (capi.ColDesc * 2)(capi.ColDesc(int(D), 1), capi.ColDesc(int(D), 1))
    pw = serialize(col(0))
    src = C.create_string_buffer(1 << 16)
    nbytes = C.c_size_t(0)
    L = capi.lib()
    rc = L.qe_debug_generate(descs, 2, (C.c_void_p * 2)(words.ctypes.data, pw.ctypes.data), (C.c_size_t * 2)(words.size, pw.size),
                             1, 1, (C.c_int32 * 1)(), 0, 0, 0, -1, 0, src, len(src), None, 0, C.byref(nbytes))
    assert rc == 0, L.qe_last_error().decode()
    body = src.value.decode().split("struct P")[1]
    assert "qe_cmp_lt(in.c0, in.c1)" in body and "(in.c0 < in.c1)" in body
    # arguments anywhere else are refused
    bad = serialize(fn(Function.ADD, col(0), col(1)))
    bad[-1] |= QE_CMP_IEEE << 8
    rc = L.qe_debug_generate(descs, 2, (C.c_void_p * 1)(bad.ctypes.data), (C.c_size_t * 1)(bad.size),
                             0, 1, (C.c_int32 * 1)(), 0, 0, 0, -1, 0, src, len(src), None, 0, C.byref(nbytes))
    assert rc != 0 and "takes no argument" in L.qe_last_error().decode()
    bad = serialize(lt)
    bad[-1] |= 2 << 8
    rc = L.qe_debug_generate(descs, 2, (C.c_void_p * 1)(bad.ctypes.data), (C.c_size_t * 1)(bad.size),
                             1, 0, (C.c_int32 * 1)(), 0, 0, 0, -1, 0, src, len(src), None, 0, C.byref(nbytes))
    assert rc != 0 and "QE_CMP_IEEE" in L.qe_last_error().decode()


def test_cubin_cache_memory_disk_and_aot(tmp_path):
    """A compiled pipeline comes from memory, from the ahead-of-time table baked into the library, from
    $QE_CACHE_DIR, or -- last -- from NVRTC (qe_cubin_cache.h).  Run in fresh processes: the caches are per process."""
    import subprocess
    import sys
    code = r"""
import ctypes as C, sys
sys.path.insert(0, %r); sys.path.insert(0, %r)
from queryengine_b200 import capi, workloads as W
from queryengine_b200.aot import compile_one
from queryengine_b200.ast import DataType, Function, col, fn
st = (C.c_double * 6)()
def stats():
    capi.check(capi.lib().qe_jit_stats(st)); return [int(x) for x in st[:4]] + [st[4], int(st[5])]
# a BASELINE pipeline: in the ahead-of-time table (any literal: values are parameters)
compile_one("aot", W.single_column_schema(), W.single_column_filter(0.123), [col(0, DataType.INT64, "col")])
s1 = stats()
# not a BASELINE pipeline: disk cache or NVRTC
D = DataType.DOUBLE
odd = fn(Function.CMP_GT, fn(Function.SUB, col(0), col(1)), 17.0)
compile_one("odd", [("a", D, True), ("b", D, False)], odd, [col(1)])
s2 = stats()
compile_one("odd again", [("a", D, True), ("b", D, False)], odd, [col(1)])
s3 = stats()
print("STATS", s1, s2, s3)
""" % (ROOT, os.path.join(ROOT, "tests"))
    env = dict(os.environ, QE_CACHE_DIR=str(tmp_path / "cache"))

    def run(extra_env=None):
        out = subprocess.run([sys.executable, "-c", code], env=dict(env, **(extra_env or {})), capture_output=True, text=True, check=True).stdout
        s1, s2, s3 = eval(out.split("STATS", 1)[1].strip().replace("] [", "], ["))
        return s1, s2, s3

    s1, s2, s3 = run()
    assert s1[5] >= 10                                         # entries of the ahead-of-time table
    assert s1[:4] == [0, 1, 0, 0]                              # BASELINE pipeline: from the table, no compiler involved
    assert s2[:4] == [0, 1, 0, 1] and s2[4] > 0                # first process: NVRTC, written to the cache directory
    assert s3[:4] == [1, 1, 0, 1]                              # again in the same process: memory
    assert len(list((tmp_path / "cache").glob("*.cubin"))) == 1
    s1, s2, s3 = run()
    assert s2[:4] == [0, 1, 1, 0] and s2[4] == 0               # second process: from disk, NVRTC never ran
    # without any compiler the ahead-of-time and disk entries still serve
    s1, s2, s3 = run({"QE_DISABLE_NVRTC": "1"})
    assert s1[:4] == [0, 1, 0, 0] and s2[:4] == [0, 1, 1, 0]


def test_type_errors_follow_the_reference_typecheck():
    # M/evaluator/TypeCheck.kt:50-94; message format of invalidTypes (:123-125)
    rc, _, _, err = _generate([(D, False), (B, False)], None, [fn(Function.ADD, col(0), col(1, B))], compile_=False)
    assert rc == capi.QE_ERR_PROGRAM and "Invalid operand types for [ADD] [DOUBLE, BOOLEAN]" in err
    rc, _, _, err = _generate([(D, False), (B, False)], None, [fn(Function.CMP_EQ, col(0), col(1, B))], compile_=False)
    assert rc == capi.QE_ERR_PROGRAM and "[CMP_EQ]" in err
    rc, _, _, err = _generate([(D, False), (B, False)], None, [fn(Function.IF, col(0), 1.0, 2.0)], compile_=False)
    assert rc == capi.QE_ERR_PROGRAM and "[IF]" in err
    rc, _, _, err = _generate([(D, False)], col(0), [], compile_=False)
    assert rc == capi.QE_ERR_PROGRAM and "BOOLEAN" in err
    rc, _, _, err = _generate([(D, False), (B, False)], None, [col(1, B)], [AggregationFunction.SUM], compile_=False)
    assert rc == capi.QE_ERR_PROGRAM
    rc, _, _, err = _generate([(D, False)], None, [col(0)], [AggregationFunction.ANY], compile_=False)
    assert rc == capi.QE_ERR_UNSUPPORTED       # TODO("ANY") in the reference, M/evaluator/Accumulators.kt:16-17
    rc, _, _, err = _generate([(D, False)], None, [col(5)], compile_=False)
    assert rc == capi.QE_ERR_PROGRAM and "out of range" in err
    # AND of two booleans is accepted (the reference's rule at TypeCheck.kt:80 is a bug)
    rc, _, _, err = _generate([(B, True), (B, True)], fn(Function.AND, col(0, B), col(1, B)), [], compile_=False)
    assert rc == 0, err


def test_malformed_programs_are_rejected():
    L = capi.lib()
    descs = (capi.ColDesc * 1)(capi.ColDesc(1, 0))

    def run(words):
        w = np.asarray(words, dtype=np.uint32)
        ptrs = (C.c_void_p * 1)(w.ctypes.data)
        lens = (C.c_size_t * 1)(w.size)
        return L.qe_debug_generate(descs, 1, ptrs, lens, 0, 1, None, 0, 0, 0, -1, 0, None, 0, None, 0, None)
    assert run([int(Function.ADD)]) == capi.QE_ERR_PROGRAM                     # stack underflow
    assert run([32, 32]) == capi.QE_ERR_PROGRAM                                # two values left
    assert run([33, 0]) == capi.QE_ERR_PROGRAM                                 # truncated constant
    assert run([99]) == capi.QE_ERR_PROGRAM                                    # unknown opcode
    assert run([32]) == capi.QE_OK


def test_no_gpu_fails_loudly():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    h = C.c_void_p()
    rc = capi.lib().qe_ctx_create(0, C.byref(h))
    assert rc == capi.QE_ERR_CUDA and b"no CPU fallback" in capi.lib().qe_last_error()
    with pytest.raises(capi.QeError):
        from queryengine_b200.engine import Context
        Context(0)


def test_agg_combine_is_fixed_order_and_matches_accumulator_rules():
    from queryengine_b200.engine import combine_aggregate
    A = AggregationFunction
    P = capi.AggPartial
    parts = [P(1.5, 2, -1.0, 3.0), P(0.0, 0, float("inf"), float("-inf")), P(2.25, 3, -0.0, 0.0)]
    assert combine_aggregate(A.SUM, parts) == 3.75
    assert combine_aggregate(A.COUNT, parts) == 5.0
    assert combine_aggregate(A.AVG, parts) == 0.75
    assert combine_aggregate(A.MIN, parts) == -1.0 and combine_aggregate(A.MAX, parts) == 3.0
    empty = [P(0.0, 0, 0.0, 0.0)] * 3
    assert combine_aggregate(A.SUM, empty) is None and combine_aggregate(A.COUNT, empty) == 0.0
    assert combine_aggregate(A.MIN, []) is None
    # order matters for floating-point sums: the combination is by array (rank) order
    big = [P(1e16, 1, 0, 0), P(1.0, 1, 0, 0), P(-1e16, 1, 0, 0)]
    assert combine_aggregate(A.SUM, big) == (1e16 + 1.0) + -1e16


def test_toolkit_nvrtc_is_used_even_when_an_older_one_is_already_loaded():
    """PyTorch bundles libnvrtc 12.8, whose ptxas rejects the 256-bit loads of sm_100.  With it
    loaded first (as in bench.py, which imports torch), the library must still pick the
    toolkit's 12.9 copy -- or fall back to 128-bit tile shapes -- and compile."""
    import glob
    import subprocess
    import sys
    import sysconfig
    site = sysconfig.get_paths()["purelib"]
    old = glob.glob(os.path.join(site, "nvidia", "cuda_nvrtc", "lib", "libnvrtc.so.12"))
    if not old:
        pytest.skip("no bundled libnvrtc to collide with")
    code = f"""
import ctypes, sys
sys.path.insert(0, {ROOT!r}); sys.path.insert(0, {os.path.join(ROOT, 'tests')!r})
ctypes.CDLL({old[0]!r}, mode=ctypes.RTLD_GLOBAL)
from test_capi_cpu import _generate, I
from queryengine_b200.ast import Function, col, fn
rc, src, nbytes, err = _generate([(I, False)], fn(Function.CMP_LT, col(0, I), 1e6), [col(0, I)], steps=4, vec=4)
assert rc == 0, err
print("ok", nbytes)
"""
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0 and "ok" in out.stdout, out.stderr[-2000:]


def test_string_dictionary_codes_in_programs_and_generated_code():
    # STRING = dictionary codes: the reference's TypeCheck lets strings into CMP_EQ / CMP_NE, IF branches,
    # bare columns and COUNT only (M/evaluator/TypeCheck.kt:65-71,86-94,118-120)
    from queryengine_b200.engine import StringDictionary
    S = DataType.STRING
    d = StringDictionary()
    assert [d.encode(x) for x in ("t", "f", "t", "")] == [0, 1, 0, 2] and d.decode(1) == "f" and len(d) == 3
    codes, valid = d.encode_column(["f", None, "zz"])
    assert list(codes) == [1, 0, 3] and list(valid) == [True, False, True]
    assert list(serialize(StringLiteralExpression("f"), d)) == [35, 1]
    # T/evaluator/CompilerTest.kt:104-120: IF(cond, "t", "f") compiles to a select between two codes
    e = FunctionExpression(Function.IF, [ColumnExpression("cond", 0, B), StringLiteralExpression("t"), StringLiteralExpression("f")], S)
    rc, src, nbytes, err = _generate([(B, True)], None, [e], dictionary=d)
    assert rc == 0, err
    # (the codes themselves are kernel parameters: lit_lo[slot])
    assert "const qe_i32 t0 = in.c0 ? ((qe_i32)p.lit_lo[0]) : ((qe_i32)p.lit_lo[1]);" in src and "qe_i32 o0;" in src and nbytes > 1000
    # string equality against a literal and between columns, in the ring filter, with a compacted STRING output
    f = fn(Function.OR, fn(Function.CMP_EQ, col(0, S), StringLiteralExpression("zz")), fn(Function.CMP_NE, col(0, S), col(1, S)))
    rc, src, nbytes, err = _generate([(S, True), (S, False)], f, [col(1, S)], dictionary=d)
    assert rc == 0, err
    assert "(in.c0 == ((qe_i32)p.lit_lo[0]))" in src and "(in.c0 != in.c1)" in src and "qe_stage1_i32" in src and "qe_filter_ring_body<P>" in src
    # what TypeCheck rejects is rejected here with its message
    for bad in (fn(Function.CMP_LT, col(0, S), col(1, S)), fn(Function.ADD, col(0, S), 1.0),
                fn(Function.CMP_EQ, col(0, S), 1.0), fn(Function.IF, lit(True), col(0, S), 1.0), fn(Function.NOT, col(0, S))):
        rc, _, _, err = _generate([(S, True), (S, False)], None, [bad], compile_=False, dictionary=d)
        assert rc == capi.QE_ERR_PROGRAM and "STRING" in err, err
    rc, _, _, err = _generate([(S, True)], None, [col(0, S)], [AggregationFunction.COUNT], dictionary=d)
    assert rc == 0, err
    rc, _, _, err = _generate([(S, True)], None, [col(0, S)], [AggregationFunction.MIN], compile_=False, dictionary=d)
    assert rc == capi.QE_ERR_PROGRAM


def test_jni_shim_matches_the_c_abi():
    # bindings/jni/qegpu_jni.c cannot be built here (no JDK): check its syntax and its use of every qe_* entry
    # point against include/qegpu.h with the stand-in jni.h
    import subprocess
    r = subprocess.run(["gcc", "-fsyntax-only", "-Wall", "-Werror", "-Wno-unused-parameter",
                        "-I", os.path.join(ROOT, "bindings", "jni", "jni_stub"), "-I", os.path.join(ROOT, "include"),
                        os.path.join(ROOT, "bindings", "jni", "qegpu_jni.c")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    kt = open(os.path.join(ROOT, "bindings", "kotlin", "net", "jhorstmann", "queryengine", "gpu", "QeGpu.kt")).read()
    c = open(os.path.join(ROOT, "bindings", "jni", "qegpu_jni.c")).read()
    declared = set(re.findall(r"external fun (\w+)\(", kt))
    defined = set(re.findall(r"QE_JNI\((\w+)\)\(", c))
    assert declared == defined and len(declared) >= 15, declared ^ defined


def test_groupby_code_of_every_variant_compiles():
    # TPC-H Q1 shape: two STRING keys, eight aggregates, a filter (GroupByAggregationOperator.kt:21-75)
    from queryengine_b200.engine import StringDictionary
    S = DataType.STRING
    A = AggregationFunction
    d = StringDictionary()
    schema = [(S, False), (S, True), (D, False), (D, False), (D, True), (D, False), (D, False)]
    rf, ls, qty, ep, disc, tax, sd = [col(i, t) for i, (t, _) in enumerate(schema)]
    net = fn(Function.MUL, ep, fn(Function.SUB, 1.0, disc))
    exprs = [fn(Function.CMP_LE, sd, 10471.0), rf, ls, qty, ep, net, fn(Function.MUL, net, fn(Function.ADD, 1.0, tax)), qty, ep, disc, rf]
    aggs = [A.SUM, A.SUM, A.SUM, A.MIN, A.AVG, A.AVG, A.MAX, A.COUNT]
    L = capi.lib()
    descs = (capi.ColDesc * len(schema))(*[capi.ColDesc(int(t), 1 if n else 0) for t, n in schema])
    words = [serialize(e, d) for e in exprs]
    ptrs = (C.c_void_p * len(words))(*[w.ctypes.data for w in words])
    lens = (C.c_size_t * len(words))(*[w.size for w in words])
    aggarr = (C.c_int32 * len(aggs))(*[int(a) for a in aggs])
    for variant, body in ((0, "qe_groupby_private_body<P>"), (2, "qe_groupby_copies_body<P>"), (1, "qe_groupby_atomic_body<P>"),
                          (3, "qe_groupby_reg_body<P>"), (4, "SCOPIES = 32"), (5, "SCOPIES = 8"), (6, "qe_groupby_shared_body<P>"),
                          (7, "SCOPIES = 16"), (8, "SCOPIES = 4"), (9, "SCOPIES = 2"), (10, "qe_groupby_scatter_body<P>"),
                          (11, "qe_groupby_gather_body<P>")):
        src = C.create_string_buffer(1 << 17)
        nb = C.c_size_t(0)
        rc = L.qe_debug_generate_groupby(descs, len(schema), ptrs, lens, 1, 2, aggarr, len(aggs), variant, 1, src, len(src), C.byref(nb))
        assert rc == 0, L.qe_last_error().decode()
        text = src.value.decode()
        assert body in text and nb.value > 10000
        # SUM(qty) and AVG(qty) share one sum word, SUM(ep) and AVG(ep) another: 1 + 2 + (net: sum + count) + (gross: min + count) + (disc: max, count shared with net? no: own) ...
        assert "NKEYS = 2" in text and "k[1] = 0ull; nulls |= 2u;" in text
        assert text.count("a[2 * WS] = ") == 1 if variant == 0 else True


def test_malformed_programs_are_rejected_not_crashed():
    # random word streams through the parser / binder (no GPU needed): every outcome is a status code
    L = capi.lib()
    rng = np.random.default_rng(99)
    schema = [(D, True), (B, False), (DataType.STRING, True), (I, False)]
    descs = (capi.ColDesc * len(schema))(*[capi.ColDesc(int(t), 1 if n else 0) for t, n in schema])
    ops = list(range(17)) + [32, 33, 34, 35, 36, 99, 255]
    seen = set()
    for trial in range(3000):
        n = int(rng.integers(1, 24))
        w = np.zeros(n, np.uint32)
        for i in range(n):
            op = ops[int(rng.integers(len(ops)))]
            arg = int(rng.integers(0, 6)) if rng.random() < 0.9 else int(rng.integers(0, 1 << 24))
            w[i] = op | (arg << 8)
        ptrs = (C.c_void_p * 1)(w.ctypes.data)
        lens = (C.c_size_t * 1)(w.size)
        aggarr = (C.c_int32 * 1)(0)
        nb = C.c_size_t(0)
        rc = L.qe_debug_generate(descs, len(schema), ptrs, lens, 0, 1, aggarr, 0, 0, 0, -1, 0, None, 0, None, 0, C.byref(nb))
        seen.add(rc)
        assert rc in (capi.QE_OK, capi.QE_ERR_PROGRAM, capi.QE_ERR_INVALID_ARGUMENT, capi.QE_ERR_UNSUPPORTED), (rc, list(w))
        rc = L.qe_debug_generate_groupby(descs, len(schema), ptrs, lens, 0, 1, aggarr, 0, int(rng.integers(3)), 0, None, 0, C.byref(nb))
        assert rc in (capi.QE_OK, capi.QE_ERR_PROGRAM, capi.QE_ERR_INVALID_ARGUMENT, capi.QE_ERR_UNSUPPORTED), (rc, list(w))
    assert capi.QE_OK in seen and capi.QE_ERR_PROGRAM in seen
