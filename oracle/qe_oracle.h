/*
 * qe_oracle.h -- TEST INFRASTRUCTURE ONLY.
 *
 * CPU restatement (plain C) of jhorstmann/queryengine's row-at-a-time hot
 * path: expression evaluation, filter, projection, memory scan, global and
 * group-by aggregation.  It exists to CHECK the CUDA path and to be timed as
 * the CPU baseline.  Nothing in queryengine_b200/ (the product) may include,
 * link or call it; only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs do.
 *
 * Parity status: the reference is Kotlin on the JVM and no JVM exists in this
 * image, so the oracle cannot be validated against the reference itself.  It
 * is pinned against every golden vector the reference's own tests hold for
 * this path (tests/test_oracle_golden.py, citing CompilerTest.kt, QueryTest.kt,
 * ByteCodeCompilerTest.kt, SimpleSumBenchmark.java).  Behaviour the reference's
 * tests do not pin (NaN / +-0 ordering, MOD, DIV by zero, MIN/MAX/AVG/COUNT)
 * is restated from the Java SE specification of the JDK primitives the
 * reference calls; for those cases parity is "unpinned by reference tests".
 *
 * Reference citations use M/ = src/main/java/net/jhorstmann/queryengine/ and
 * T/ = src/test/java/net/jhorstmann/queryengine/ under /root/reference.
 */
#ifndef QE_ORACLE_H
#define QE_ORACLE_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

/* M/data/Schema.kt:3-5 -- DataType { STRING, DOUBLE, BOOLEAN } (ordinals kept) */
enum { ORC_STRING = 0, ORC_DOUBLE = 1, ORC_BOOLEAN = 2 };

/* M/ast/Functions.kt:7-22 -- Function enum, ordinals kept */
enum {
    ORC_AND = 0, ORC_OR, ORC_IF, ORC_NOT, ORC_UNARY_MINUS, ORC_UNARY_PLUS,
    ORC_MUL, ORC_DIV, ORC_MOD, ORC_ADD, ORC_SUB,
    ORC_CMP_LT, ORC_CMP_LE, ORC_CMP_GE, ORC_CMP_GT, ORC_CMP_EQ, ORC_CMP_NE,
    ORC_FUNCTION_COUNT
};

/* M/ast/Functions.kt:24-26 -- AggregationFunction enum, ordinals kept */
enum { ORC_MIN = 0, ORC_MAX, ORC_SUM, ORC_COUNT, ORC_AVG, ORC_ANY, ORC_ALL };

/* M/evaluator/Compiler.kt:5-7 -- Mode.  The three back ends disagree on a few
 * edge cases; the oracle reproduces each one so the
 * disagreement itself is testable.  ORC_MODE_BYTECODE is the graded semantics
 * (Double.compare total order, NOT null -> null, unary plus = identity). */
enum { ORC_MODE_INTERPRETER = 0, ORC_MODE_CLOSURE = 1, ORC_MODE_BYTECODE = 2 };

/* Boxed value: the JVM's java.lang.Double / Boolean / String.  A NULL pointer
 * is the Kotlin null. */
typedef struct orc_value {
    int tag;                /* ORC_STRING / ORC_DOUBLE / ORC_BOOLEAN, or ORC_INT for COUNT */
    union { double d; int b; const char *s; int32_t i; } u;
} orc_value;
enum { ORC_INT = 3 };       /* CountAccumulator.finish() returns a Kotlin Int, M/evaluator/Accumulators.kt:33-35 */

typedef struct orc_expr orc_expr;
typedef struct orc_table orc_table;
typedef struct orc_op orc_op;

/* ---- expression trees (M/ast/Expressions.kt:6-62) ---- */
orc_expr *orc_column(int index, int dtype);            /* ColumnExpression(name,index,dataType) */
orc_expr *orc_numeric(double v);                       /* NumericLiteralExpression */
orc_expr *orc_boolean(int v);                          /* BooleanLiteralExpression */
orc_expr *orc_string(const char *s);                   /* StringLiteralExpression (copied) */
orc_expr *orc_function(int fn, int nops, orc_expr *const *ops); /* FunctionExpression; takes ownership of ops */
void      orc_expr_free(orc_expr *e);

/* RowCallable.invoke (M/evaluator/Compiler.kt:9-11).  Returns a boxed value
 * valid until the next orc_eval on this thread, or NULL for null.  *err is set
 * to 1 where the reference would throw (e.g. NOT null in INTERPRETER mode). */
const orc_value *orc_eval(const orc_expr *e, const orc_value *const *row, int mode, int *err);

/* ---- MemoryTable (M/data/MemoryTable.kt:7-19): row-major, boxed ---- */
orc_table *orc_table_create(int ncols, const int *dtypes, int64_t nrows);
/* validity: Arrow-style LSB-first bitmap, 1 = non-null; NULL = all non-null */
int  orc_table_set_double(orc_table *t, int col, const double *values, const uint8_t *validity);
int  orc_table_set_boolean(orc_table *t, int col, const uint8_t *values01, const uint8_t *validity);
int  orc_table_set_string(orc_table *t, int col, const char *const *values /* NULL entry = null */);
void orc_table_free(orc_table *t);

/* ---- operators (the files under M/operator/) ---- */
orc_op *orc_memory_source(const orc_table *t, const int *indices, int nidx);         /* MemorySourceOperator.kt */
orc_op *orc_memory_source_range(const orc_table *t, const int *indices, int nidx,
                                int64_t row0, int64_t row1);                         /* same, rows [row0,row1) */
orc_op *orc_filter(orc_op *source, const orc_expr *predicate, int mode);             /* FilterOperator.kt */
orc_op *orc_projection(orc_op *source, const orc_expr *const *exprs, int n, int mode); /* ProjectionOperator.kt */
orc_op *orc_global_aggregation(orc_op *source, const int *fns, int n);               /* GlobalAggregationOperator.kt */
orc_op *orc_groupby_aggregation(orc_op *source, int group_count, const int *fns, int n); /* GroupByAggregationOperator.kt */
void    orc_open(orc_op *op);
const orc_value *const *orc_next(orc_op *op);   /* NULL at end; row valid until next call */
void    orc_close(orc_op *op);
int     orc_width(const orc_op *op);            /* number of columns in rows this operator returns */
int     orc_error(const orc_op *op);            /* 1 if the reference would have thrown */
void    orc_op_free(orc_op *op);                /* frees the whole chain below it too */

/* Operator.forEach (M/operator/Operators.kt:13-23) into columnar buffers, for
 * comparison with the GPU results.  out_values[c] points to `capacity` doubles
 * (booleans stored as 0.0/1.0, Int as its value; strings not supported here),
 * out_valid[c] to `capacity` bytes (1 = non-null).  Returns the row count, or
 * -1 on error / overflow. */
int64_t orc_drain(orc_op *op, int ncols, double *const *out_values, uint8_t *const *out_valid,
                  int64_t capacity);
/* Same pull loop, but only counts rows and folds values into a checksum so the
 * compiler cannot drop the work; used for timing. */
int64_t orc_drain_count(orc_op *op, double *checksum);

/* ---- timed baseline: the same operator chain, row-range sharded over
 * `threads` pthreads (the reference itself is single-threaded).  Plan =
 * source -> [filter] -> [projection] -> [global aggregation].  Writes
 * aggregate results (value, is_null) when naggs > 0, otherwise the number of
 * rows that reached the sink.  Returns elapsed seconds. */
double orc_timed_pipeline(const orc_table *t, const int *indices, int nidx,
                          const orc_expr *filter,
                          const orc_expr *const *proj, int nproj,
                          const int *agg_fns, int naggs,
                          int mode, int threads,
                          double *agg_values, int *agg_is_null, int64_t *rows_out);

#ifdef __cplusplus
}
#endif
#endif
