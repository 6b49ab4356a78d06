/*
 * qe_oracle.c -- TEST INFRASTRUCTURE ONLY (see qe_oracle.h).
 *
 * Row-at-a-time, boxed-value restatement of the reference's evaluator and
 * operators.  Every function cites the reference lines it follows.  Boxes
 * created during evaluation come from a bump arena (the JVM's TLAB analogue)
 * that the scan resets whenever it hands out the next row.
 */
#define _GNU_SOURCE
#include "qe_oracle.h"

#include <math.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

/* ------------------------------------------------------------------ arena */

typedef struct orc_arena {
    char  *base;
    size_t cap;
    size_t used;
} orc_arena;

static void arena_init(orc_arena *a, size_t cap)
{
    a->base = (char *)malloc(cap);
    a->cap = cap;
    a->used = 0;
}

static void *arena_alloc(orc_arena *a, size_t n)
{
    n = (n + 15u) & ~(size_t)15u;
    if (a->used + n > a->cap) {
        /* expression trees are tiny; a row never needs more than a few KB.
         * Growing would invalidate live boxes, so fail hard instead. */
        abort();
    }
    void *p = a->base + a->used;
    a->used += n;
    return p;
}

static orc_value *box_double(orc_arena *a, double d)
{
    orc_value *v = (orc_value *)arena_alloc(a, sizeof *v);
    v->tag = ORC_DOUBLE;
    v->u.d = d;
    return v;
}

static orc_value *box_boolean(orc_arena *a, int b)
{
    orc_value *v = (orc_value *)arena_alloc(a, sizeof *v);
    v->tag = ORC_BOOLEAN;
    v->u.b = b ? 1 : 0;
    return v;
}

/* ------------------------------------------------- JDK primitive restatements */

static int64_t raw_bits(double d)
{
    int64_t b;
    memcpy(&b, &d, sizeof b);
    return b;
}

/* java.lang.Double.doubleToLongBits: every NaN collapses to 0x7ff8000000000000 */
static int64_t double_to_long_bits(double d)
{
    if (d != d) return (int64_t)0x7ff8000000000000LL;
    return raw_bits(d);
}

/* java.lang.Double.compare(double,double) -- Java SE spec: -0.0 < 0.0, NaN is
 * equal to itself and greater than +Infinity.  Called by the reference at
 * M/evaluator/BytecodeCompiler.kt:300 and (through Double.compareTo) at
 * M/evaluator/Interpreter.kt:104-107. */
static int jdk_double_compare(double a, double b)
{
    if (a < b) return -1;
    if (a > b) return 1;
    int64_t x = double_to_long_bits(a);
    int64_t y = double_to_long_bits(b);
    return x == y ? 0 : (x < y ? -1 : 1);
}

/* java.lang.Math.min / max (M/evaluator/Accumulators.kt:62,80) */
static double jdk_math_min(double a, double b)
{
    if (a != a) return a;
    if (a == 0.0 && b == 0.0 && raw_bits(b) == (int64_t)0x8000000000000000ULL) return b;
    return a <= b ? a : b;
}

static double jdk_math_max(double a, double b)
{
    if (a != a) return a;
    if (a == 0.0 && b == 0.0 && raw_bits(a) == (int64_t)0x8000000000000000ULL) return b;
    return a >= b ? a : b;
}

/* Kotlin `a == b` on Any -> a.equals(b): Double.equals compares
 * doubleToLongBits (NaN == NaN, 0.0 != -0.0); M/evaluator/Interpreter.kt:102-103,
 * M/evaluator/ClosureCompiler.kt:125-126. */
static int boxed_equals(const orc_value *a, const orc_value *b)
{
    if (a->tag != b->tag) return 0;
    switch (a->tag) {
    case ORC_DOUBLE:  return double_to_long_bits(a->u.d) == double_to_long_bits(b->u.d);
    case ORC_BOOLEAN: return a->u.b == b->u.b;
    case ORC_STRING:  return strcmp(a->u.s, b->u.s) == 0;
    case ORC_INT:     return a->u.i == b->u.i;
    }
    return 0;
}

/* Comparable.compareTo for the three boxed types (M/evaluator/Interpreter.kt:104-107,
 * M/evaluator/BytecodeCompiler.kt:299-307) */
static int boxed_compare(const orc_value *a, const orc_value *b)
{
    switch (a->tag) {
    case ORC_DOUBLE:  return jdk_double_compare(a->u.d, b->u.d);
    case ORC_BOOLEAN: return (a->u.b > b->u.b) - (a->u.b < b->u.b);   /* Boolean.compare */
    case ORC_STRING: { int c = strcmp(a->u.s, b->u.s); return (c > 0) - (c < 0); }
    }
    return 0;
}

/* ------------------------------------------------------------ expressions */

enum { EX_COLUMN, EX_NUMERIC, EX_BOOLEAN, EX_STRING, EX_FUNCTION };

struct orc_expr {
    int kind;
    int index;          /* EX_COLUMN */
    int dtype;          /* EX_COLUMN */
    int fn;             /* EX_FUNCTION */
    int nops;
    orc_expr **ops;
    orc_value literal;  /* pre-boxed literal: the JVM boxes it once at parse time */
    char *str;
};

static orc_expr *expr_new(int kind)
{
    orc_expr *e = (orc_expr *)calloc(1, sizeof *e);
    e->kind = kind;
    return e;
}

orc_expr *orc_column(int index, int dtype)
{
    orc_expr *e = expr_new(EX_COLUMN);
    e->index = index;
    e->dtype = dtype;
    return e;
}

orc_expr *orc_numeric(double v)
{
    orc_expr *e = expr_new(EX_NUMERIC);
    e->literal.tag = ORC_DOUBLE;
    e->literal.u.d = v;
    return e;
}

orc_expr *orc_boolean(int v)
{
    orc_expr *e = expr_new(EX_BOOLEAN);
    e->literal.tag = ORC_BOOLEAN;
    e->literal.u.b = v ? 1 : 0;
    return e;
}

orc_expr *orc_string(const char *s)
{
    orc_expr *e = expr_new(EX_STRING);
    e->str = strdup(s);
    e->literal.tag = ORC_STRING;
    e->literal.u.s = e->str;
    return e;
}

orc_expr *orc_function(int fn, int nops, orc_expr *const *ops)
{
    orc_expr *e = expr_new(EX_FUNCTION);
    e->fn = fn;
    e->nops = nops;
    e->ops = (orc_expr **)malloc(sizeof(orc_expr *) * (size_t)(nops > 0 ? nops : 1));
    for (int i = 0; i < nops; i++) e->ops[i] = ops[i];
    return e;
}

void orc_expr_free(orc_expr *e)
{
    if (!e) return;
    for (int i = 0; i < e->nops; i++) orc_expr_free(e->ops[i]);
    free(e->ops);
    free(e->str);
    free(e);
}

/* FunctionType of each Function (M/ast/Functions.kt:7-22): 0 LOGIC, 1 ARITHMETIC, 2 COMPARISON */
static int function_is_logic(int fn)
{
    return fn == ORC_AND || fn == ORC_OR || fn == ORC_IF || fn == ORC_NOT;
}

typedef struct eval_ctx {
    orc_arena *arena;
    const orc_value *const *row;
    int mode;
    int err;
} eval_ctx;

static const orc_value *eval(eval_ctx *c, const orc_expr *e);

/* Interpreter.visitFunction, M/evaluator/Interpreter.kt:29-109.
 * Mode differences:
 *   CLOSURE: ordering comparisons are IEEE on unboxed doubles
 *            (M/evaluator/ClosureCompiler.kt:127-130); NOT null -> null (:115).
 *   BYTECODE: Double.compare (M/evaluator/BytecodeCompiler.kt:300); NOT null ->
 *            null (:346-351); UNARY_PLUS restated as identity (the emitted
 *            bytecode at :341 is invalid, so there is nothing to reproduce).
 *   INTERPRETER: NOT null throws (cast of null to Boolean, Interpreter.kt:92). */
static const orc_value *eval_function(eval_ctx *c, const orc_expr *e)
{
    const orc_value *args[3] = {0, 0, 0};
    int fn = e->fn;

    if (!function_is_logic(fn)) {
        /* Interpreter.kt:35-42: evaluate every operand, then null-check */
        int any_null = 0;
        for (int i = 0; i < e->nops; i++) {
            args[i] = eval(c, e->ops[i]);
            if (!args[i]) any_null = 1;
        }
        if (any_null) return NULL;
    }

    switch (fn) {
    case ORC_IF: {                                       /* Interpreter.kt:46-53 */
        const orc_value *cond = eval(c, e->ops[0]);
        if (!cond) return NULL;
        return cond->u.b ? eval(c, e->ops[1]) : eval(c, e->ops[2]);
    }
    case ORC_AND: {                                      /* Interpreter.kt:54-72 */
        const orc_value *p = eval(c, e->ops[0]);
        if (!p) {
            const orc_value *q = eval(c, e->ops[1]);
            if (!q) return NULL;
            if (!q->u.b) return box_boolean(c->arena, 0);
            return NULL;
        }
        if (p->u.b) return eval(c, e->ops[1]);
        return box_boolean(c->arena, 0);
    }
    case ORC_OR: {                                       /* Interpreter.kt:73-91 */
        const orc_value *p = eval(c, e->ops[0]);
        if (!p) {
            const orc_value *q = eval(c, e->ops[1]);
            if (!q) return NULL;
            if (q->u.b) return box_boolean(c->arena, 1);
            return NULL;
        }
        if (p->u.b) return box_boolean(c->arena, 1);
        return eval(c, e->ops[1]);
    }
    case ORC_NOT: {                                      /* Interpreter.kt:92 */
        const orc_value *p = eval(c, e->ops[0]);
        if (!p) {
            if (c->mode == ORC_MODE_INTERPRETER) c->err = 1;   /* NullPointerException / TypeCastException */
            return NULL;
        }
        return box_boolean(c->arena, !p->u.b);
    }
    case ORC_UNARY_PLUS:  return args[0];                                        /* :94 */
    case ORC_UNARY_MINUS: return box_double(c->arena, -args[0]->u.d);            /* :95 */
    case ORC_ADD: return box_double(c->arena, args[0]->u.d + args[1]->u.d);      /* :96 */
    case ORC_SUB: return box_double(c->arena, args[0]->u.d - args[1]->u.d);      /* :97 */
    case ORC_MUL: return box_double(c->arena, args[0]->u.d * args[1]->u.d);      /* :98 */
    case ORC_DIV: return box_double(c->arena, args[0]->u.d / args[1]->u.d);      /* :99 */
    case ORC_MOD: return box_double(c->arena, fmod(args[0]->u.d, args[1]->u.d)); /* :100, JVM DREM == C fmod */
    case ORC_CMP_EQ: return box_boolean(c->arena, boxed_equals(args[0], args[1]));   /* :102 */
    case ORC_CMP_NE: return box_boolean(c->arena, !boxed_equals(args[0], args[1]));  /* :103 */
    case ORC_CMP_LT: case ORC_CMP_LE: case ORC_CMP_GE: case ORC_CMP_GT: {
        int r;
        if (c->mode == ORC_MODE_CLOSURE) {
            /* ClosureCompiler.kt:127-130: (a as Double) < (b as Double), IEEE */
            if (args[0]->tag != ORC_DOUBLE || args[1]->tag != ORC_DOUBLE) { c->err = 1; return NULL; }
            double a = args[0]->u.d, b = args[1]->u.d;
            r = fn == ORC_CMP_LT ? a < b : fn == ORC_CMP_LE ? a <= b : fn == ORC_CMP_GE ? a >= b : a > b;
        } else {
            int cmp = boxed_compare(args[0], args[1]);   /* Interpreter.kt:104-107 */
            r = fn == ORC_CMP_LT ? cmp < 0 : fn == ORC_CMP_LE ? cmp <= 0 : fn == ORC_CMP_GE ? cmp >= 0 : cmp > 0;
        }
        return box_boolean(c->arena, r);
    }
    }
    c->err = 1;
    return NULL;
}

static const orc_value *eval(eval_ctx *c, const orc_expr *e)
{
    switch (e->kind) {
    case EX_COLUMN:   return c->row[e->index];           /* Interpreter.kt:25-27 */
    case EX_NUMERIC:                                     /* Interpreter.kt:13-15 */
    case EX_BOOLEAN:                                     /* :17-19 */
    case EX_STRING:   return &e->literal;                /* :21-23 */
    case EX_FUNCTION: return eval_function(c, e);
    }
    c->err = 1;
    return NULL;
}

static __thread orc_arena tls_arena;

const orc_value *orc_eval(const orc_expr *e, const orc_value *const *row, int mode, int *err)
{
    if (!tls_arena.base) arena_init(&tls_arena, 1u << 16);
    tls_arena.used = 0;
    eval_ctx c = { &tls_arena, row, mode, 0 };
    const orc_value *r = eval(&c, e);
    if (err) *err = c.err;
    return r;
}

/* ------------------------------------------------------------------ table */

/* MemoryTable(schema, values: List<List<Any?>>), M/data/MemoryTable.kt:7.
 * Row-major array of pointers to individually boxed values, like the JVM heap
 * layout the reference scans (three dependent loads per value). */
struct orc_table {
    int ncols;
    int *dtypes;
    int64_t nrows;
    const orc_value **cells;   /* nrows * ncols pointers, row-major */
    orc_value **col_boxes;     /* per column: backing array of boxes */
    char ***col_strings;       /* per column: owned string copies */
};

orc_table *orc_table_create(int ncols, const int *dtypes, int64_t nrows)
{
    orc_table *t = (orc_table *)calloc(1, sizeof *t);
    t->ncols = ncols;
    t->nrows = nrows;
    t->dtypes = (int *)malloc(sizeof(int) * (size_t)ncols);
    memcpy(t->dtypes, dtypes, sizeof(int) * (size_t)ncols);
    t->cells = (const orc_value **)calloc((size_t)(nrows * ncols) + 1, sizeof(orc_value *));
    t->col_boxes = (orc_value **)calloc((size_t)ncols, sizeof(orc_value *));
    t->col_strings = (char ***)calloc((size_t)ncols, sizeof(char **));
    return t;
}

static int bit_is_set(const uint8_t *bitmap, int64_t i)
{
    return !bitmap || ((bitmap[i >> 3] >> (i & 7)) & 1);
}

int orc_table_set_double(orc_table *t, int col, const double *values, const uint8_t *validity)
{
    if (col < 0 || col >= t->ncols || t->dtypes[col] != ORC_DOUBLE) return -1;
    free(t->col_boxes[col]);
    orc_value *boxes = (orc_value *)malloc(sizeof(orc_value) * (size_t)(t->nrows + 1));
    t->col_boxes[col] = boxes;
    for (int64_t i = 0; i < t->nrows; i++) {
        boxes[i].tag = ORC_DOUBLE;
        boxes[i].u.d = values[i];
        t->cells[i * t->ncols + col] = bit_is_set(validity, i) ? &boxes[i] : NULL;
    }
    return 0;
}

int orc_table_set_boolean(orc_table *t, int col, const uint8_t *values01, const uint8_t *validity)
{
    if (col < 0 || col >= t->ncols || t->dtypes[col] != ORC_BOOLEAN) return -1;
    free(t->col_boxes[col]);
    orc_value *boxes = (orc_value *)malloc(sizeof(orc_value) * (size_t)(t->nrows + 1));
    t->col_boxes[col] = boxes;
    for (int64_t i = 0; i < t->nrows; i++) {
        boxes[i].tag = ORC_BOOLEAN;
        boxes[i].u.b = values01[i] ? 1 : 0;
        t->cells[i * t->ncols + col] = bit_is_set(validity, i) ? &boxes[i] : NULL;
    }
    return 0;
}

int orc_table_set_string(orc_table *t, int col, const char *const *values)
{
    if (col < 0 || col >= t->ncols || t->dtypes[col] != ORC_STRING) return -1;
    orc_value *boxes = (orc_value *)malloc(sizeof(orc_value) * (size_t)(t->nrows + 1));
    char **strs = (char **)calloc((size_t)t->nrows + 1, sizeof(char *));
    t->col_boxes[col] = boxes;
    t->col_strings[col] = strs;
    for (int64_t i = 0; i < t->nrows; i++) {
        if (values[i]) {
            strs[i] = strdup(values[i]);
            boxes[i].tag = ORC_STRING;
            boxes[i].u.s = strs[i];
            t->cells[i * t->ncols + col] = &boxes[i];
        } else {
            t->cells[i * t->ncols + col] = NULL;
        }
    }
    return 0;
}

void orc_table_free(orc_table *t)
{
    if (!t) return;
    for (int c = 0; c < t->ncols; c++) {
        free(t->col_boxes[c]);
        if (t->col_strings[c]) {
            for (int64_t i = 0; i < t->nrows; i++) free(t->col_strings[c][i]);
            free(t->col_strings[c]);
        }
    }
    free(t->col_boxes);
    free(t->col_strings);
    free(t->cells);
    free(t->dtypes);
    free(t);
}

/* ------------------------------------------------------------ accumulators */

/* M/evaluator/Accumulators.kt:21-107.  `count` is a Kotlin Int: 32-bit, wraps. */
typedef struct orc_acc {
    int fn;
    int32_t count;
    double sum;      /* SUM / AVG */
    double minmax;   /* MIN / MAX */
} orc_acc;

static void acc_init(orc_acc *a, int fn)            /* initAccumulator, Accumulators.kt:9-19 */
{
    a->fn = fn;
    a->count = 0;
    a->sum = 0.0;
    a->minmax = fn == ORC_MIN ? INFINITY : -INFINITY;    /* :57, :75 */
}

static void acc_accumulate(orc_acc *a, const orc_value *v)
{
    a->count = (int32_t)((uint32_t)a->count + 1u);       /* Int overflow wraps */
    switch (a->fn) {
    case ORC_COUNT: break;                                                  /* :29-31 */
    case ORC_SUM: case ORC_AVG: a->sum += v->u.d; break;                    /* :42-45, :96-99 */
    case ORC_MIN: a->minmax = jdk_math_min(a->minmax, v->u.d); break;       /* :60-63 */
    case ORC_MAX: a->minmax = jdk_math_max(a->minmax, v->u.d); break;       /* :78-81 */
    }
}

/* finish(): returns 1 and fills *out when non-null, 0 for null */
static int acc_finish(const orc_acc *a, orc_value *out)
{
    switch (a->fn) {
    case ORC_COUNT: out->tag = ORC_INT; out->u.i = a->count; return 1;      /* :33-35 */
    case ORC_SUM:                                                           /* :47-53 */
        if (a->count == 0) return 0;
        out->tag = ORC_DOUBLE; out->u.d = a->sum;
        return 1;
    case ORC_MIN: case ORC_MAX:                                             /* :65-71, :83-89 */
        if (a->count == 0) return 0;
        out->tag = ORC_DOUBLE; out->u.d = a->minmax;
        return 1;
    case ORC_AVG:                                                           /* :101-107 */
        if (a->count == 0) return 0;
        out->tag = ORC_DOUBLE; out->u.d = a->sum / a->count;
        return 1;
    }
    return 0;
}

/* -------------------------------------------------------------- operators */

enum { OP_SOURCE, OP_FILTER, OP_PROJECTION, OP_GLOBAL_AGG, OP_GROUPBY_AGG };

typedef struct group_entry {
    const orc_value **key;     /* group_count boxed values (owned copies), NULL = null */
    orc_acc *acc;
    struct group_entry *next_in_bucket;
} group_entry;

struct orc_op {
    int kind;
    int width;
    int err;
    int mode;
    orc_op *source;
    orc_arena *arena;          /* shared down the chain, owned by the scan */

    /* OP_SOURCE: MemorySourceOperator.kt:5-36 */
    const orc_table *table;
    int *indices;
    int64_t idx, begin, end;
    const orc_value **rowbuf;  /* the single shared row buffer, :8 */

    /* OP_FILTER */
    const orc_expr *predicate;

    /* OP_PROJECTION */
    const orc_expr **exprs;

    /* OP_GLOBAL_AGG / OP_GROUPBY_AGG */
    int *fns;
    int nfns;
    int group_count;
    const orc_value **result;  /* GlobalAggregationOperator.kt:8 */
    orc_value *result_boxes;
    int result_ready;
    group_entry **entries;     /* insertion order (LinkedHashMap) */
    int64_t nentries, cap_entries, iter;
    group_entry **buckets;
    int64_t nbuckets;
};

static orc_op *op_new(int kind)
{
    orc_op *op = (orc_op *)calloc(1, sizeof *op);
    op->kind = kind;
    return op;
}

orc_op *orc_memory_source_range(const orc_table *t, const int *indices, int nidx, int64_t row0, int64_t row1)
{
    orc_op *op = op_new(OP_SOURCE);
    op->table = t;
    op->width = nidx;
    op->indices = (int *)malloc(sizeof(int) * (size_t)(nidx > 0 ? nidx : 1));
    memcpy(op->indices, indices, sizeof(int) * (size_t)nidx);
    op->rowbuf = (const orc_value **)calloc((size_t)nidx + 1, sizeof(orc_value *));
    op->begin = row0;
    op->end = row1;
    op->arena = (orc_arena *)malloc(sizeof(orc_arena));
    arena_init(op->arena, 1u << 16);
    return op;
}

orc_op *orc_memory_source(const orc_table *t, const int *indices, int nidx)
{
    return orc_memory_source_range(t, indices, nidx, 0, t->nrows);
}

orc_op *orc_filter(orc_op *source, const orc_expr *predicate, int mode)
{
    orc_op *op = op_new(OP_FILTER);
    op->source = source;
    op->arena = source->arena;
    op->predicate = predicate;
    op->mode = mode;
    op->width = source->width;
    return op;
}

orc_op *orc_projection(orc_op *source, const orc_expr *const *exprs, int n, int mode)
{
    orc_op *op = op_new(OP_PROJECTION);
    op->source = source;
    op->arena = source->arena;
    op->mode = mode;
    op->width = n;
    op->exprs = (const orc_expr **)malloc(sizeof(orc_expr *) * (size_t)(n > 0 ? n : 1));
    memcpy(op->exprs, exprs, sizeof(orc_expr *) * (size_t)n);
    return op;
}

static orc_op *agg_new(int kind, orc_op *source, int group_count, const int *fns, int n)
{
    orc_op *op = op_new(kind);
    op->source = source;
    op->arena = source->arena;
    op->group_count = group_count;
    op->nfns = n;
    op->width = group_count + n;
    op->fns = (int *)malloc(sizeof(int) * (size_t)(n > 0 ? n : 1));
    memcpy(op->fns, fns, sizeof(int) * (size_t)n);
    op->result = (const orc_value **)calloc((size_t)op->width + 1, sizeof(orc_value *));
    op->result_boxes = (orc_value *)calloc((size_t)op->width + 1, sizeof(orc_value));
    return op;
}

orc_op *orc_global_aggregation(orc_op *source, const int *fns, int n)
{
    return agg_new(OP_GLOBAL_AGG, source, 0, fns, n);
}

orc_op *orc_groupby_aggregation(orc_op *source, int group_count, const int *fns, int n)
{
    return agg_new(OP_GROUPBY_AGG, source, group_count, fns, n);
}

int orc_width(const orc_op *op) { return op->width; }

int orc_error(const orc_op *op)
{
    for (; op; op = op->source) if (op->err) return 1;
    return 0;
}

/* Arrays.hashCode / contentEquals over the boxed key (GroupByAggregationOperator.kt:8-16).
 * Only equality semantics matter for the result; any consistent hash will do. */
static uint64_t key_hash(const orc_value *const *key, int n)
{
    uint64_t h = 1469598103934665603ULL;
    for (int i = 0; i < n; i++) {
        uint64_t x = 0x9e3779b97f4a7c15ULL;
        const orc_value *v = key[i];
        if (v) {
            switch (v->tag) {
            case ORC_DOUBLE:  x = (uint64_t)double_to_long_bits(v->u.d); break;
            case ORC_BOOLEAN: x = (uint64_t)v->u.b + 17; break;
            case ORC_STRING:  for (const char *s = v->u.s; *s; s++) x = x * 131 + (uint8_t)*s; break;
            }
        }
        h = (h ^ x) * 1099511628211ULL;
    }
    return h;
}

static int key_equals(const orc_value *const *a, const orc_value *const *b, int n)
{
    for (int i = 0; i < n; i++) {
        if (!a[i] || !b[i]) { if (a[i] != b[i] && (a[i] || b[i])) return 0; continue; }
        if (!boxed_equals(a[i], b[i])) return 0;
    }
    return 1;
}

static const orc_value *value_clone(const orc_value *v)
{
    if (!v) return NULL;
    orc_value *c = (orc_value *)malloc(sizeof *c);
    *c = *v;
    if (v->tag == ORC_STRING) c->u.s = strdup(v->u.s);
    return c;
}

static void groups_clear(orc_op *op)
{
    for (int64_t i = 0; i < op->nentries; i++) {
        group_entry *g = op->entries[i];
        for (int k = 0; k < op->group_count; k++) {
            if (g->key[k]) {
                if (g->key[k]->tag == ORC_STRING) free((void *)g->key[k]->u.s);
                free((void *)g->key[k]);
            }
        }
        free(g->key);
        free(g->acc);
        free(g);
    }
    free(op->entries);
    free(op->buckets);
    op->entries = NULL;
    op->buckets = NULL;
    op->nentries = op->cap_entries = op->nbuckets = 0;
}

static group_entry *groups_get(orc_op *op, const orc_value *const *row)
{
    if (!op->buckets) {
        op->nbuckets = 1024;
        op->buckets = (group_entry **)calloc((size_t)op->nbuckets, sizeof(group_entry *));
    }
    uint64_t h = key_hash(row, op->group_count);
    group_entry *g = op->buckets[h % (uint64_t)op->nbuckets];
    for (; g; g = g->next_in_bucket)
        if (key_equals(g->key, row, op->group_count)) return g;
    /* computeIfAbsent, GroupByAggregationOperator.kt:35-37 */
    g = (group_entry *)calloc(1, sizeof *g);
    g->key = (const orc_value **)calloc((size_t)op->group_count + 1, sizeof(orc_value *));
    for (int k = 0; k < op->group_count; k++) g->key[k] = value_clone(row[k]);
    g->acc = (orc_acc *)calloc((size_t)op->nfns + 1, sizeof(orc_acc));
    for (int i = 0; i < op->nfns; i++) acc_init(&g->acc[i], op->fns[i]);
    g->next_in_bucket = op->buckets[h % (uint64_t)op->nbuckets];
    op->buckets[h % (uint64_t)op->nbuckets] = g;
    if (op->nentries == op->cap_entries) {
        op->cap_entries = op->cap_entries ? op->cap_entries * 2 : 64;
        op->entries = (group_entry **)realloc(op->entries, sizeof(group_entry *) * (size_t)op->cap_entries);
    }
    op->entries[op->nentries++] = g;
    return g;
}

void orc_open(orc_op *op)
{
    switch (op->kind) {
    case OP_SOURCE:                                   /* MemorySourceOperator.kt:10-12 */
        op->idx = op->begin;
        break;
    case OP_FILTER:                                   /* FilterOperator.kt:6-8 */
    case OP_PROJECTION:                               /* ProjectionOperator.kt:6-8 */
        orc_open(op->source);
        break;
    case OP_GLOBAL_AGG: {                             /* GlobalAggregationOperator.kt:10-25 */
        orc_acc acc[64];
        for (int i = 0; i < op->nfns; i++) acc_init(&acc[i], op->fns[i]);
        /* source.forEach: open, pull until null, close (Operators.kt:13-23) */
        orc_open(op->source);
        const orc_value *const *row;
        while ((row = orc_next(op->source)) != NULL) {
            for (int i = 0; i < op->nfns; i++) {
                const orc_value *v = row[i];
                if (v) acc_accumulate(&acc[i], v);    /* :17-20 nulls are skipped */
            }
        }
        orc_close(op->source);
        for (int i = 0; i < op->nfns; i++)            /* :24 */
            op->result[i] = acc_finish(&acc[i], &op->result_boxes[i]) ? &op->result_boxes[i] : NULL;
        op->result_ready = 1;
        break;
    }
    case OP_GROUPBY_AGG: {                            /* GroupByAggregationOperator.kt:21-49 */
        groups_clear(op);
        orc_open(op->source);
        const orc_value *const *row;
        if (op->source->width != op->group_count + op->nfns) op->err = 1;   /* :29-31 */
        while (!op->err && (row = orc_next(op->source)) != NULL) {
            group_entry *g = groups_get(op, row);
            for (int i = 0; i < op->nfns; i++) {
                const orc_value *v = row[op->group_count + i];
                if (v) acc_accumulate(&g->acc[i], v);
            }
        }
        orc_close(op->source);
        op->iter = 0;
        op->result_ready = 1;
        break;
    }
    }
}

const orc_value *const *orc_next(orc_op *op)
{
    switch (op->kind) {
    case OP_SOURCE: {                                 /* MemorySourceOperator.kt:18-36 */
        int64_t i = op->idx;
        if (i < op->end) {
            op->idx = i + 1;
            op->arena->used = 0;                      /* previous row's temporaries are garbage now */
            const orc_value **row = op->table->cells + i * op->table->ncols;
            for (int j = 0; j < op->width; j++) op->rowbuf[j] = row[op->indices[j]];
            return op->rowbuf;
        }
        return NULL;
    }
    case OP_FILTER: {                                 /* FilterOperator.kt:14-25 */
        for (;;) {
            const orc_value *const *row = orc_next(op->source);
            if (!row) return NULL;
            eval_ctx c = { op->arena, row, op->mode, 0 };
            const orc_value *res = eval(&c, op->predicate);
            if (c.err) op->err = 1;
            if (res != NULL && res->u.b) return row;  /* :20 null predicate drops the row */
        }
    }
    case OP_PROJECTION: {                             /* ProjectionOperator.kt:15-19 */
        const orc_value *const *row = orc_next(op->source);
        if (!row) return NULL;
        const orc_value **out = (const orc_value **)arena_alloc(op->arena, sizeof(orc_value *) * (size_t)(op->width + 1));
        for (int k = 0; k < op->width; k++) {
            eval_ctx c = { op->arena, row, op->mode, 0 };
            out[k] = eval(&c, op->exprs[k]);
            if (c.err) op->err = 1;
        }
        return out;
    }
    case OP_GLOBAL_AGG: {                             /* GlobalAggregationOperator.kt:32-36 */
        if (!op->result_ready) return NULL;
        op->result_ready = 0;
        return op->result;
    }
    case OP_GROUPBY_AGG: {                            /* GroupByAggregationOperator.kt:56-75 */
        if (!op->result_ready) { op->err = 1; return NULL; }      /* "Operator not opened" */
        if (op->iter >= op->nentries) return NULL;
        group_entry *g = op->entries[op->iter++];
        for (int k = 0; k < op->group_count; k++) op->result[k] = g->key[k];
        for (int i = 0; i < op->nfns; i++) {
            int k = op->group_count + i;
            op->result[k] = acc_finish(&g->acc[i], &op->result_boxes[k]) ? &op->result_boxes[k] : NULL;
        }
        return op->result;
    }
    }
    return NULL;
}

void orc_close(orc_op *op)
{
    switch (op->kind) {
    case OP_SOURCE:                                   /* MemorySourceOperator.kt:14-16 */
        for (int j = 0; j < op->width; j++) op->rowbuf[j] = NULL;
        break;
    case OP_FILTER: case OP_PROJECTION:
        orc_close(op->source);
        break;
    case OP_GLOBAL_AGG:                               /* GlobalAggregationOperator.kt:28-30 */
        op->result_ready = 0;
        break;
    case OP_GROUPBY_AGG:                              /* GroupByAggregationOperator.kt:51-54 */
        groups_clear(op);
        op->result_ready = 0;
        break;
    }
}

void orc_op_free(orc_op *op)
{
    if (!op) return;
    orc_op_free(op->source);
    if (op->kind == OP_SOURCE) {
        free(op->arena->base);
        free(op->arena);
    }
    if (op->kind == OP_GROUPBY_AGG) groups_clear(op);
    free(op->indices);
    free(op->rowbuf);
    free(op->exprs);
    free(op->fns);
    free(op->result);
    free(op->result_boxes);
    free(op);
}

/* ----------------------------------------------------------- pull helpers */

static double value_as_double(const orc_value *v)
{
    switch (v->tag) {
    case ORC_DOUBLE:  return v->u.d;
    case ORC_BOOLEAN: return v->u.b ? 1.0 : 0.0;
    case ORC_INT:     return (double)v->u.i;
    }
    return 0.0;
}

/* Operator.forEach / mapTo (M/operator/Operators.kt:13-32) into columns */
int64_t orc_drain(orc_op *op, int ncols, double *const *out_values, uint8_t *const *out_valid, int64_t capacity)
{
    int64_t n = 0;
    orc_open(op);
    const orc_value *const *row;
    while ((row = orc_next(op)) != NULL) {
        if (n >= capacity) { orc_close(op); return -1; }
        for (int c = 0; c < ncols; c++) {
            const orc_value *v = row[c];
            out_valid[c][n] = v ? 1 : 0;
            out_values[c][n] = v ? value_as_double(v) : 0.0;
        }
        n++;
    }
    orc_close(op);
    return orc_error(op) ? -1 : n;
}

int64_t orc_drain_count(orc_op *op, double *checksum)
{
    int64_t n = 0;
    double acc = 0.0;
    int w = op->width;
    orc_open(op);
    const orc_value *const *row;
    while ((row = orc_next(op)) != NULL) {
        for (int c = 0; c < w; c++)
            if (row[c] && row[c]->tag != ORC_STRING) acc += value_as_double(row[c]);
        n++;
    }
    orc_close(op);
    if (checksum) *checksum = acc;
    return n;
}

/* --------------------------------------------------- timed CPU baseline */

typedef struct shard {
    const orc_table *t;
    const int *indices; int nidx;
    const orc_expr *filter;
    const orc_expr *const *proj; int nproj;
    const int *agg_fns; int naggs;
    int mode;
    int64_t row0, row1;
    /* results */
    int64_t rows;
    double checksum;
    orc_acc acc[64];
} shard;

/* One thread = one independent copy of the reference's pull chain over a row
 * range (the reference is single-threaded; sharding is the baseline's own). */
static void *shard_run(void *arg)
{
    shard *s = (shard *)arg;
    orc_op *op = orc_memory_source_range(s->t, s->indices, s->nidx, s->row0, s->row1);
    if (s->filter) op = orc_filter(op, s->filter, s->mode);
    if (s->nproj > 0) op = orc_projection(op, s->proj, s->nproj, s->mode);
    if (s->naggs > 0) {
        /* GlobalAggregationOperator.open's loop, kept open-coded so the
         * per-shard accumulators can be merged afterwards */
        for (int i = 0; i < s->naggs; i++) acc_init(&s->acc[i], s->agg_fns[i]);
        orc_open(op);
        const orc_value *const *row;
        int64_t n = 0;
        while ((row = orc_next(op)) != NULL) {
            for (int i = 0; i < s->naggs; i++)
                if (row[i]) acc_accumulate(&s->acc[i], row[i]);
            n++;
        }
        orc_close(op);
        s->rows = n;
    } else {
        s->rows = orc_drain_count(op, &s->checksum);
    }
    orc_op_free(op);
    return NULL;
}

double orc_timed_pipeline(const orc_table *t, const int *indices, int nidx,
                          const orc_expr *filter,
                          const orc_expr *const *proj, int nproj,
                          const int *agg_fns, int naggs,
                          int mode, int threads,
                          double *agg_values, int *agg_is_null, int64_t *rows_out)
{
    if (threads < 1) threads = 1;
    if (naggs > 64) return -1.0;
    shard *sh = (shard *)calloc((size_t)threads, sizeof *sh);
    pthread_t *tid = (pthread_t *)calloc((size_t)threads, sizeof *tid);
    int64_t per = (t->nrows + threads - 1) / threads;
    struct timespec t0, t1;
    clock_gettime(CLOCK_MONOTONIC, &t0);
    for (int i = 0; i < threads; i++) {
        shard *s = &sh[i];
        s->t = t; s->indices = indices; s->nidx = nidx; s->filter = filter;
        s->proj = proj; s->nproj = nproj; s->agg_fns = agg_fns; s->naggs = naggs; s->mode = mode;
        s->row0 = per * i < t->nrows ? per * i : t->nrows;
        s->row1 = per * (i + 1) < t->nrows ? per * (i + 1) : t->nrows;
        if (threads == 1) shard_run(s);
        else pthread_create(&tid[i], NULL, shard_run, s);
    }
    if (threads > 1) for (int i = 0; i < threads; i++) pthread_join(tid[i], NULL);
    clock_gettime(CLOCK_MONOTONIC, &t1);

    int64_t rows = 0;
    for (int i = 0; i < threads; i++) rows += sh[i].rows;
    if (rows_out) *rows_out = rows;
    /* merge shard accumulators in shard order */
    for (int k = 0; k < naggs; k++) {
        orc_acc m;
        acc_init(&m, agg_fns[k]);
        for (int i = 0; i < threads; i++) {
            const orc_acc *a = &sh[i].acc[k];
            if (a->count == 0) continue;
            m.count = (int32_t)((uint32_t)m.count + (uint32_t)a->count);
            m.sum += a->sum;
            if (agg_fns[k] == ORC_MIN) m.minmax = jdk_math_min(m.minmax, a->minmax);
            if (agg_fns[k] == ORC_MAX) m.minmax = jdk_math_max(m.minmax, a->minmax);
        }
        orc_value v;
        int nonnull = acc_finish(&m, &v);
        if (agg_is_null) agg_is_null[k] = !nonnull;
        if (agg_values) agg_values[k] = nonnull ? value_as_double(&v) : 0.0;
    }
    free(sh);
    free(tid);
    return (double)(t1.tv_sec - t0.tv_sec) + 1e-9 * (double)(t1.tv_nsec - t0.tv_nsec);
}
