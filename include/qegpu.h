/*
 * qegpu.h -- C ABI of libqegpu.so, the B200 (sm_100a) implementation of
 * jhorstmann/queryengine's filter/scan hot path.
 *
 * The reference has no FFI, plugin or SPI of any kind; its seams are three
 * abstract classes and one enum (M/ = src/main/java/net/jhorstmann/queryengine/):
 *
 *   Operator    { open(); next(): Array<Any?>?; close() }   M/operator/Operators.kt:5-11
 *   RowCallable { invoke(row: Array<Any?>): Any? }          M/evaluator/Compiler.kt:9-11
 *   Table       { schema; getScanOperator(projection) }     M/data/Table.kt:6-10
 *   Mode        { INTERPRETER, CLOSURE_COMPILER, BYTECODE_COMPILER }   M/evaluator/Compiler.kt:5-7
 *
 * and the one place they are assembled, buildPhysicalPlan (M/evaluator/Planner.kt:30-63).
 * This library replaces the subtree that function builds for
 *     Scan -> [Filter] -> [Projection] -> [GlobalAggregation]
 * with one call.  Each entry point below names the reference interface it
 * stands in for.  INTEGRATION.md shows the Kotlin/JNI binding a maintainer
 * would add (Mode.GPU + GpuPipelineOperator).
 *
 * Conventions
 *   - every function returns a qe_status (0 = OK); the message of the last
 *     failure on the calling thread is qe_last_error().  No C++ exception
 *     crosses this boundary.  (The reference throws unchecked exceptions.)
 *   - handles are opaque, created and destroyed by the library.  A handle is
 *     used by one thread at a time; different handles may be used concurrently.
 *   - host buffers passed in or filled are owned by the caller and are not
 *     retained after the call returns.
 *   - there is no CPU fallback: without a CUDA device qe_ctx_create fails.
 */
#ifndef QEGPU_H
#define QEGPU_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define QE_API __attribute__((visibility("default")))
#else
#define QE_API
#endif

typedef enum qe_status {
    QE_OK = 0,
    QE_ERR_INVALID_ARGUMENT = 1,
    QE_ERR_PROGRAM = 2,        /* malformed program or operand types the reference's TypeCheck rejects */
    QE_ERR_CUDA = 3,
    QE_ERR_UNSUPPORTED = 4,    /* e.g. ANY / ALL aggregates, more columns than QE_MAX_PIPELINE_COLUMNS */
    QE_ERR_CAPACITY = 5,       /* more rows passed the filter than out_capacity_rows */
    QE_ERR_JIT = 6,
    QE_ERR_STATE = 7
} qe_status;

/* Physical column types.  STRING, DOUBLE and BOOLEAN are the reference's DataType
 * members (M/data/Schema.kt:3-5, same ordinals); INT64 and DATE32 are
 * extensions that widen exactly to DOUBLE when loaded.  CONTRACT: every value of an
 * INT64 column satisfies |x| < 2^53 -- the column stands for the DOUBLE values the
 * reference would hold, and only below 2^53 is that one-to-one.  The library does not
 * scan uploads (qe_table_generate rejects ranges beyond it); with larger values the
 * generated code (exact integer compare against literals) and the interpreter kernels
 * (compare after widening) may disagree, and neither is "the reference's answer".
 * STRING columns hold dictionary codes: the caller interns every string of a
 * context (column values AND literals) in ONE dictionary, so that equal codes
 * <=> equal strings.  That is all the reference's type checker lets a query do
 * with strings: CMP_EQ / CMP_NE, IF branches, bare columns, COUNT and group keys
 * (M/evaluator/TypeCheck.kt:65-71,86-94,118-120). */
typedef enum qe_type {
    QE_TYPE_STRING = 0,        /* int32_t[n] dictionary codes */
    QE_TYPE_DOUBLE = 1,        /* double[n] */
    QE_TYPE_BOOLEAN = 2,       /* bitmap, LSB first */
    QE_TYPE_INT64 = 3,         /* int64_t[n] */
    QE_TYPE_DATE32 = 4         /* int32_t[n], days since epoch */
} qe_type;

/* AggregationFunction ordinals, M/ast/Functions.kt:24-26 */
typedef enum qe_agg_fn {
    QE_AGG_FN_MIN = 0, QE_AGG_FN_MAX = 1, QE_AGG_FN_SUM = 2, QE_AGG_FN_COUNT = 3, QE_AGG_FN_AVG = 4
} qe_agg_fn;

/* ---- flattened expression program -----------------------------------------
 * A tree (M/ast/Expressions.kt:6-62) in postfix order, one 32-bit word per
 * node:  word = opcode | (argument << 8).
 *   opcodes 0..16   the reference's Function enum ordinals (M/ast/Functions.kt:7-22):
 *                   AND OR IF NOT UNARY_MINUS UNARY_PLUS MUL DIV MOD ADD SUB
 *                   CMP_LT CMP_LE CMP_GE CMP_GT CMP_EQ CMP_NE; operands were
 *                   pushed left to right (IF: condition, then, else).  Their argument
 *                   is 0, except that CMP_LT / LE / GE / GT take QE_CMP_IEEE (below).
 *   QE_OP_LOAD_COL  ColumnExpression; argument = column index in the qe_table
 *   QE_OP_CONST_F64 NumericLiteralExpression; the next two words are the low
 *                   and high halves of the IEEE-754 bits
 *   QE_OP_CONST_BOOL BooleanLiteralExpression; argument = 0 / 1
 *   QE_OP_CONST_STR StringLiteralExpression; the next word is its dictionary code (int32)
 */
enum {
    QE_OP_LOAD_COL = 32,
    QE_OP_CONST_F64 = 33,
    QE_OP_CONST_BOOL = 34,
    QE_OP_CONST_STR = 35
};
/* Argument of CMP_LT / CMP_LE / CMP_GE / CMP_GT.  0: java.lang.Double.compare's total order (-0.0 < 0.0, NaN equal to itself
 * and above +Infinity) -- Mode.INTERPRETER and Mode.BYTECODE_COMPILER, M/evaluator/Interpreter.kt:104-107,
 * BytecodeCompiler.kt:300.  QE_CMP_IEEE: the unboxed `(a as Double) < (b as Double)` of Mode.CLOSURE_COMPILER,
 * M/evaluator/ClosureCompiler.kt:127-130 -- any ordering comparison with a NaN is false, -0.0 equals 0.0.  It is the one
 * place where that mode's results differ from the other two on well-typed trees; CMP_EQ / CMP_NE are
 * boxed equals in all three (ClosureCompiler.kt:125-126). */
#define QE_CMP_IEEE 1u
#define QE_MAX_PROGRAM_WORDS 1024
#define QE_MAX_STACK 32
#define QE_MAX_PIPELINE_COLUMNS 16
#define QE_MAX_PROJECTIONS 8

typedef struct qe_ctx qe_ctx;
typedef struct qe_program qe_program;
typedef struct qe_table qe_table;
typedef struct qe_result qe_result;

typedef struct qe_col_desc {
    int32_t type;              /* qe_type */
    int32_t nullable;          /* 1: a validity bitmap is stored */
} qe_col_desc;

/* Partial aggregate of one partition, for combining across GPUs in fixed rank order. */
typedef struct qe_agg_partial {
    double  sum;
    int64_t count;             /* non-null inputs seen */
    double  min;
    double  max;
} qe_agg_partial;

typedef struct qe_pipeline_desc {
    const qe_program *filter;          /* FilterOperator predicate, may be NULL */
    const qe_program *const *proj;     /* ProjectionOperator expressions */
    int32_t nproj;
    const int32_t *agg_fns;            /* GlobalAggregationOperator functions; agg k consumes projection k */
    int32_t nagg;                      /* 0, or == nproj */
    int64_t row_begin;                 /* batch = rows [row_begin, row_end); row_begin % 128 == 0 */
    int64_t row_end;                   /* -1 = to the end of the table */
    int64_t out_capacity_rows;         /* rows the outputs may hold; 0 = worst case (all rows of this batch) */
    uint32_t flags;                    /* QE_RUN_* */
    int64_t sel_offset;                /* added to the table row index in the selection vector (0 normally; a batch that
                                          holds rows [g, g+n) of a larger table at rows [s, s+n) of this one passes g - s);
                                          row + sel_offset must fit 32 bits */
} qe_pipeline_desc;

enum {
    QE_RUN_NO_SELECTION = 1u << 0,     /* do not materialise the selection vector */
    QE_RUN_INTERPRETER  = 1u << 1,     /* use the generic interpreter kernels instead of generated code */
    QE_RUN_APPEND       = 1u << 2,     /* *result holds earlier batches of the same pipeline: append this batch's rows
                                          (order preserved; capacity is the first run's out_capacity_rows) */
    QE_RUN_NO_VALUES    = 1u << 3      /* filter pipelines: the selection vector is the whole output -- the projections are
                                          neither evaluated nor written.  For callers that hold the table themselves
                                          (the reference's MemoryTable is in host memory, M/data/MemoryTable.kt:7, and
                                          FilterOperator returns the very row it was given, M/operator/FilterOperator.kt:20-22) */
};

/* ---- context: one CUDA device, one stream, a device memory pool ------------ */
QE_API int qe_ctx_create(int device, qe_ctx **out);
QE_API int qe_ctx_destroy(qe_ctx *ctx);
QE_API int qe_ctx_synchronize(qe_ctx *ctx);
/* cudaStream_t all work of this context is enqueued on (for event timing by the caller) */
QE_API void *qe_ctx_stream(qe_ctx *ctx);
/* options: "jit" (1 default / 0 = always interpret), "steps", "vec" (tile shape overrides, 0 = auto),
 * "grid_per_sm" (0 = auto), "staging" (filter kernel variant: -1 auto = 2; 0 register-sourced; 1 bulk-copy stages,
 * look-back in line; 2 warp-specialised ring of bulk-copy stages with deferred scatter), "stages" (0 = auto, 2..32),
 * ring variant: "stage_kb" (bytes of column data per stage, 0 = auto), "tile_rows" (0 = auto), "ncw" (consumer
 * warps, 0 = auto: 20 for a single 8-byte column, else 16), "defer" (tiles between evaluation and scatter, 0 = auto);
 * group-by: "groupby_registers" (1 default: accumulators in registers when the cost rule allows; 0 never; 2 whenever
 * the registers suffice), "groupby_lanes" (1 default; 0 skips the 32-copies-per-CTA variant), "groupby_twopass" (1 default: more groups than a CTA's shared memory holds are partitioned by key hash and aggregated per partition before global atomics are tried; 0 skips that), "groupby_warps" (0 default; 1 tries the
 * one-copy-per-warp variant, which measured slower than the single shared copy), "jit_cache_entries" (kernels kept loaded per context, default 128) */
QE_API int qe_ctx_set_option(qe_ctx *ctx, const char *key, int64_t value);
/* kernels launched by this context so far */
QE_API int64_t qe_ctx_launch_count(const qe_ctx *ctx);
/* give the device memory the context's pool holds but does not use back to the driver */
QE_API int qe_ctx_trim(qe_ctx *ctx);

/* HBM yardsticks with the pipelines' own access shapes: kind 0 reads `bytes`, 1 copies `bytes` (read + write), 2 writes
 * `bytes`; best device time of `reps` passes.  bench.py's roofline denominators for a launch's read / write mix. */
QE_API int qe_yardstick(qe_ctx *ctx, int kind, int64_t bytes, int reps, float *best_ms);

/* Where compiled pipelines came from, process-wide (stands in for the compile-once-per-expression of
 * M/evaluator/BytecodeCompiler.kt:134-175): lookups served from [0] memory, [1] the ahead-of-time table baked into
 * the library, [2] the on-disk cache ($QE_CACHE_DIR, else $XDG_CACHE_HOME/qegpu, else $HOME/.cache/qegpu; "off"
 * disables it), [3] compiled by NVRTC; [4] = milliseconds spent in NVRTC, [5] = entries of the ahead-of-time table. */
QE_API int qe_jit_stats(double out[6]);

/* pinned host memory for staging-free uploads / downloads */
QE_API int qe_host_alloc(size_t bytes, void **out);
QE_API int qe_host_free(void *p);

/* ---- program: stands in for compileExpression(expr, mode): RowCallable
 *      (M/evaluator/Compiler.kt:20-26).  Structure is checked here; operand types are
 *      checked against the table when the pipeline runs (the rules of
 *      M/evaluator/TypeCheck.kt:38-96 with the evident intent for AND/OR). */
QE_API int qe_program_create(qe_ctx *ctx, const uint32_t *words, size_t nwords, qe_program **out);
QE_API int qe_program_destroy(qe_program *program);

/* ---- table: columnar stand-in for MemoryTable (M/data/MemoryTable.kt:7-19) -- */
QE_API int qe_table_create(qe_ctx *ctx, int ncols, const qe_col_desc *cols, int64_t nrows, qe_table **out);
/* Copy rows [row0, row0+n) of one column from host memory (pinned or pageable)
 * into HBM, asynchronously on the context's stream.  `values`: see qe_type.
 * `validity`: Arrow bitmap for those n rows (bit i = row0+i non-null) or NULL
 * = all non-null.  row0 must be a multiple of 8 when a bitmap is involved. */
QE_API int qe_table_upload(qe_table *table, int col, const void *values, const uint8_t *validity,
                           int64_t row0, int64_t n);
/* Device addresses (values: which = 0, validity bitmap: which = 1) for callers
 * that produce columns on the device. */
QE_API void *qe_table_device_ptr(qe_table *table, int col, int which);
/* Synthetic column generators (counter-based, a pure function of the global
 * row index `row_offset + i`, so results do not depend on the GPU count).
 * kind: see QE_GEN_* ; p0, p1 are kind-specific; null_rate in [0,1] applies
 * to nullable columns. */
enum {
    QE_GEN_UNIFORM_U40 = 0,    /* splitmix64(seed ^ i) >> 24, as the column's type */
    QE_GEN_UNIFORM_INT = 1,    /* p0 + (hash % (p1 - p0 + 1)), integers in [p0, p1] */
    QE_GEN_UNIFORM_REAL = 2,   /* p0 + u * (p1 - p0), u in [0,1) with 53 random bits */
    QE_GEN_CENTS = 3,          /* uniform integer in [p0, p1] divided by 100.0 */
    QE_GEN_ROW_INDEX = 4       /* the global row index itself */
};
QE_API int qe_table_generate(qe_table *table, int col, int kind, uint64_t seed, double p0, double p1,
                             double null_rate, int64_t row_offset);
QE_API int qe_table_destroy(qe_table *table);

/* ---- pipeline: stands in for the operator subtree buildPhysicalPlan builds
 *      (M/evaluator/Planner.kt:30-63) and for pulling it dry (Operator.forEach,
 *      M/operator/Operators.kt:13-23).  *result may hold a result of an earlier
 *      run of the same shape, whose buffers are then reused. */
QE_API int qe_pipeline_run(qe_ctx *ctx, const qe_table *table,
                           const qe_program *filter,
                           const qe_program *const *proj, int nproj,
                           const int32_t *agg_fns, int nagg,
                           qe_result **result);
QE_API int qe_pipeline_run_ex(qe_ctx *ctx, const qe_table *table, const qe_pipeline_desc *desc,
                              qe_result **result);

/* ---- streaming: host columns in, host results out, chunk by chunk -----------
 * Stands in for the same subtree when the table starts in host memory (the reference's MemoryTable
 * always does, M/data/MemoryTable.kt:7) or is larger than the device (MemorySourceOperator scans a table of
 * any size, M/operator/MemorySourceOperator.kt:18-36): uploads, kernels and downloads of neighbouring
 * row-range chunks overlap on three streams.  Host buffers should be pinned (qe_host_alloc) for the copies
 * to be asynchronous.  Returns when every copy has completed.
 *
 * Two modes:
 *   resident (total_rows == 0)   `table` holds the whole table; chunk c is uploaded to its own rows and every
 *       chunk appends to the same device-side result (QE_RUN_APPEND), which stays readable through qe_result_*
 *       as after qe_pipeline_run.
 *   bounded (total_rows > 0)     `table` is a window of table.nrows / chunk_rows (>= 2) chunk slots used round
 *       robin; total_rows may exceed both the device memory and 2^32.  Every chunk writes into one of as many
 *       output buffer sets (worst case: a chunk's rows each), which is downloaded to its place in the caller's
 *       host buffers and reused.  Afterwards qe_result_num_rows / qe_result_agg* / qe_result_kernel_stats work
 *       on results[q]; the rows themselves are in the host buffers only (qe_result_column: QE_ERR_STATE).
 *       pipelines[q].out_capacity_rows = rows host_selection[q] / host_out_values[q] can hold (0 = total_rows);
 *       QE_ERR_CAPACITY when a filter keeps more.
 * Selection vectors hold global row indices; with QE_STREAM_CHUNK_RELATIVE (required beyond 2^32 rows) they
 * hold the index inside the chunk, and chunk_kept says which output positions belong to which chunk:
 * global row of output i = chunk * chunk_rows + sel[i] for chunk_kept[q][chunk-1] <= i < chunk_kept[q][chunk].
 * Aggregations: every chunk contributes its super-tile partials to one list, folded in row order at the end --
 * with chunk_rows a multiple of QE_AGG_SUPER_ROWS the result is bit-identical to one run over the resident
 * table (read it with qe_result_agg / qe_result_agg_partial). */
typedef struct qe_stream_gen {           /* synthetic source of one column, as qe_table_generate */
    int32_t kind;                        /* QE_GEN_*, or -1: not generated */
    uint64_t seed;
    double p0, p1, null_rate;
} qe_stream_gen;
typedef struct qe_stream_desc {
    int64_t chunk_rows;                          /* rows per chunk, a multiple of 128; 0 = 32 Mi */
    const void *const *host_values;              /* [table columns] source of each column (as qe_table_upload), NULL entry = already resident */
    const uint8_t *const *host_validity;         /* [table columns] Arrow bitmaps, or NULL */
    int32_t npipelines;
    const qe_pipeline_desc *pipelines;           /* row_begin / row_end / sel_offset / QE_RUN_APPEND are set by the library */
    qe_result **results;                         /* [npipelines] in / out, as for qe_pipeline_run_ex */
    uint32_t *const *host_selection;             /* [npipelines] destination of the selection vectors; NULL or NULL entries = not wanted */
    void *const *const *host_out_values;         /* [npipelines][nproj] destinations of the output columns */
    uint8_t *const *const *host_out_validity;    /* [npipelines][nproj] destinations of the validity bitmaps, or NULL */
    int64_t total_rows;                          /* 0: resident mode; > 0: bounded mode, rows of the whole (host) table */
    uint32_t flags;                              /* QE_STREAM_* */
    const qe_stream_gen *gen;                    /* [table columns] or NULL: columns generated on the device chunk by chunk
                                                    from the global row index gen_row_offset + row (benchmarks of tables no
                                                    host could hold) */
    int64_t gen_row_offset;
    int64_t *const *chunk_kept;                  /* [npipelines][chunks] out, or NULL: rows kept up to and including each chunk */
} qe_stream_desc;
enum {
    QE_STREAM_CHUNK_RELATIVE = 1u << 0           /* selection vectors count from the first row of their chunk */
};
QE_API int qe_pipeline_stream(qe_ctx *ctx, qe_table *table, const qe_stream_desc *desc);

/* ---- group-by: stands in for Scan -> [Filter] -> Projection -> GroupByAggregationOperator
 *      (M/operator/GroupByAggregationOperator.kt:21-75; the plan RewriteAggregates builds,
 *      M/evaluator/RewriteAggregates.kt:18-49: every non-aggregate select expression is a key).
 *      A row's key is the tuple of its `keys` values; keys compare like the reference's boxed
 *      values (Double.equals: NaN == NaN, -0.0 != 0.0; null is a key value like any other; STRING
 *      by dictionary code).  Accumulator k of a group takes `inputs[k]`, nulls skipped.  The
 *      result has one row per group, in order of first appearance (LinkedHashMap), with the
 *      nkeys key columns first and then the nagg aggregates (DOUBLE; NULL as finish() would
 *      return null); read it with qe_result_num_rows / qe_result_column.  The call is
 *      synchronous.  Few groups (what fits a CTA's shared memory) accumulate in
 *      per-thread private copies and are combined in a fixed order, so results are reproducible;
 *      more groups share one copy per CTA (shared-memory atomics), still more use global atomics
 *      (SUM / AVG then depend on the order of additions, within the tolerance stated for SUM).
 *      Needs NVRTC. */
typedef struct qe_groupby_desc {
    const qe_program *filter;          /* may be NULL */
    const qe_program *const *keys;     /* group key expressions */
    int32_t nkeys;                     /* 0..8 */
    const qe_program *const *inputs;   /* aggregate input expressions */
    const int32_t *agg_fns;            /* qe_agg_fn of each input */
    int32_t nagg;                      /* nkeys + nagg <= 16 */
    int64_t row_begin;                 /* rows [row_begin, row_end); row_begin % 128 == 0 */
    int64_t row_end;                   /* -1 = to the end of the table */
    int64_t max_groups;                /* capacity of the group table; 0 = 65536.  QE_ERR_CAPACITY when exceeded */
    uint32_t flags;                    /* QE_GROUPBY_* */
} qe_groupby_desc;
enum {
    QE_GROUPBY_ATOMIC = 1u << 0,       /* go straight to global atomics */
    QE_GROUPBY_SHARED = 1u << 1,       /* skip the private-accumulator variants: start with the shared-memory copies */
    QE_GROUPBY_TWOPASS = 1u << 2       /* go straight to the two-pass variant (rows partitioned by key hash), whatever the batch size */
};
QE_API int qe_groupby_run(qe_ctx *ctx, const qe_table *table, const qe_groupby_desc *desc, qe_result **result);
/* per-group partial aggregates and first rows, for combining row-range partitions by key on the host */
QE_API int qe_result_group_partial(qe_result *result, int64_t group, int k, qe_agg_partial *out);
QE_API int qe_result_group_first_rows(qe_result *result, int64_t *dst);
/* 3: accumulators in registers (few groups; reproducible), 0: private accumulators in shared memory (reproducible),
 * 4: one shared-memory copy per lane index (32 per CTA), 5: sixteen or eight copies per CTA, 6: one copy per warp updated
 * without atomics (reproducible), 2: four, two or one shared-memory copies per CTA, 7: two passes (rows partitioned by
 * key hash, one CTA per partition), 1: global atomics, -1: not a group-by result */
QE_API int qe_result_group_variant(const qe_result *result);
/* copies of the accumulator words per CTA of variants 4 / 5 / 2 (32, 16, 8, 4, 2 or 1); 0 for the other variants */
QE_API int qe_result_group_copies(const qe_result *result);
/* threads per CTA of the variant that produced the result (private: 512 when the groups fit, else 256) */
QE_API int qe_result_group_threads(const qe_result *result);

/* ---- results ------------------------------------------------------------- */
/* rows the pipeline emitted (kept rows; 1 for an aggregation).  Waits for the run. */
QE_API int qe_result_num_rows(qe_result *result, int64_t *nrows);
QE_API int qe_result_num_columns(const qe_result *result, int *ncols);
QE_API int qe_result_column_type(const qe_result *result, int k, int *type, int *has_validity);
/* Copy output column k to host: values as for qe_type (BOOLEAN: bitmap of
 * ceil(n/8) bytes), validity bitmap likewise (all ones when the column cannot
 * be null).  Either pointer may be NULL. */
QE_API int qe_result_column(qe_result *result, int k, void *dst_values, uint8_t *dst_validity);
/* ascending table row indices of the rows that passed the filter */
QE_API int qe_result_selection(qe_result *result, uint32_t *dst);
/* finish() of accumulator k (M/evaluator/Accumulators.kt:21-107); COUNT is returned as a double */
QE_API int qe_result_agg(qe_result *result, int k, double *value, int *is_null);
QE_API int qe_result_agg_partial(qe_result *result, int k, qe_agg_partial *out);
/* The partials of aggregate k per super-tile of QE_AGG_SUPER_ROWS rows, in row order (at most `capacity` are written,
 * *n receives how many there are).  A sum is formed in a fixed shape over the ROW INDEX -- rows of a 16 384-row
 * sub-tile in a fixed order, 64 sub-tiles per super-tile in a fixed tree, super-tiles one by one in ascending order --
 * so qe_agg_combine over the concatenated lists of all partitions gives the same bits for any number of partitions
 * or streamed chunks whose boundaries are multiples of QE_AGG_SUPER_ROWS. */
#define QE_AGG_SUPER_ROWS 1048576
QE_API int qe_result_agg_partials(qe_result *result, int k, qe_agg_partial *dst, int64_t capacity, int64_t *n);
/* device addresses of the result buffers: which = 0 values, 1 validity, 2 selection vector (k ignored) */
QE_API void *qe_result_device_ptr(qe_result *result, int k, int which);
/* device time of the main kernel of the last run that produced this result, in milliseconds */
QE_API int qe_result_kernel_ms(qe_result *result, float *ms);
/* device time of the main kernel summed over every completed run into this result, and the run count */
QE_API int qe_result_kernel_stats(qe_result *result, double *total_ms, int64_t *runs);
/* development aid: per-CTA phase cycle counters recorded when option "debug_flags" has bit 16 ([cta][8] words) */
QE_API int qe_result_debug_counters(qe_result *result, uint64_t *dst, size_t nwords);
/* 1 if the last run used generated (JIT) code, 0 for the interpreter kernels */
QE_API int qe_result_used_jit(const qe_result *result);
QE_API int qe_result_destroy(qe_result *result);

/* Combine per-partition partials in array order (deterministic) and finish. */
QE_API int qe_agg_combine(int agg_fn, const qe_agg_partial *partials, int n, double *value, int *is_null);

/* ---- several GPUs, one process ------------------------------------------------
 * The reference is one process that builds one plan (M/Main.kt:11-19, M/evaluator/Planner.kt:30); a JVM adapter
 * cannot call torch.distributed.  A qe_mgpu owns one context per device and one worker thread per context; a table is
 * split into contiguous row ranges (partition boundaries are multiples of QE_AGG_SUPER_ROWS, so SUM / AVG do not depend
 * on the number of devices), every call fans out to all devices at once and joins.  Filter / projection outputs stay
 * on their device ("gathered per partition": concatenating the partitions in order is the reference's scan order,
 * M/operator/FilterOperator.kt:17-22); the only exchange is the partial aggregates, which every device leaves in
 * pinned host memory and the calling thread folds in partition order (GlobalAggregationOperator.kt:10-25,
 * Accumulators.kt:38-54) -- for one process that is cheaper than a collective, see DESIGN.md.  Group-by results are
 * merged by key on the host (fixed-width records, partials in partition order, groups ordered by first row).
 * Programs are host objects: one created with any context of the group (qe_mgpu_ctx) runs on all of them.  STRING
 * columns of all partitions must use ONE dictionary. */
typedef struct qe_mgpu qe_mgpu;
typedef struct qe_mtable qe_mtable;
typedef struct qe_mresult qe_mresult;

QE_API int qe_mgpu_create(const int *devices, int ndevices, qe_mgpu **out);
QE_API int qe_mgpu_destroy(qe_mgpu *m);
QE_API int qe_mgpu_num_devices(const qe_mgpu *m);
QE_API qe_ctx *qe_mgpu_ctx(qe_mgpu *m, int part);
QE_API int qe_mgpu_set_option(qe_mgpu *m, const char *key, int64_t value);      /* on every context */
/* rows [*begin, *end) of partition `part` of an nrows-row table over `nparts` partitions */
QE_API int qe_partition_rows(int64_t nrows, int nparts, int part, int64_t *begin, int64_t *end);

QE_API int qe_mtable_create(qe_mgpu *m, int ncols, const qe_col_desc *cols, int64_t nrows, qe_mtable **out);
/* whole column from host memory; every device copies its row range, all at the same time */
QE_API int qe_mtable_upload(qe_mtable *t, int col, const void *values, const uint8_t *validity);
/* synthetic column from the global row index (as qe_table_generate with row_offset = the partition's first row) */
QE_API int qe_mtable_generate(qe_mtable *t, int col, int kind, uint64_t seed, double p0, double p1, double null_rate);
QE_API qe_table *qe_mtable_part(qe_mtable *t, int part);
QE_API int qe_mtable_destroy(qe_mtable *t);

/* desc->row_begin / row_end must be 0 / -1 (the whole table); out_capacity_rows applies to every partition */
QE_API int qe_mgpu_pipeline_run(qe_mgpu *m, const qe_mtable *t, const qe_pipeline_desc *desc, qe_mresult **result);
QE_API int qe_mgpu_groupby_run(qe_mgpu *m, const qe_mtable *t, const qe_groupby_desc *desc, qe_mresult **result);
/* descs[part] describes the stream of partition `part` (host pointers already offset to its first row);
 * qe_pipeline_stream runs on every device at the same time */
QE_API int qe_mgpu_pipeline_stream(qe_mgpu *m, qe_mtable *t, const qe_stream_desc *descs);

/* Merge the groups of several row-range partitions by key (what qe_mgpu_groupby_run does after the devices have
 * finished; also for callers that gathered the partitions' groups some other way, e.g. one process per GPU and an
 * all-gather of these arrays).  Keys are compared by their value bits (as qe_result_column returns them, widened to
 * 64 bits; NaN keys are canonical) and null flags; partials of one group are folded in partition order; groups come
 * out ordered by their smallest first_row -- the order one scan of the whole table would have met them in.
 * out_* arrays hold `capacity` groups: out_keys / out_key_null [g][nkeys], out_values / out_null [g][nagg]. */
typedef struct qe_group_part {
    int64_t ngroups;
    const uint64_t *keys;              /* [ngroups][nkeys] value bits */
    const uint8_t *key_null;           /* [ngroups][nkeys] 1 = the key value is null */
    const int64_t *first_row;          /* [ngroups] GLOBAL row of the group's first appearance */
    const qe_agg_partial *partials;    /* [ngroups][nagg] */
} qe_group_part;
QE_API int qe_groupby_merge(const qe_group_part *parts, int nparts, int nkeys, const int32_t *agg_fns, int nagg,
                            int64_t capacity, uint64_t *out_keys, uint8_t *out_key_null, int64_t *out_first_row,
                            double *out_values, uint8_t *out_null, int64_t *out_ngroups);

QE_API int qe_mresult_num_rows(qe_mresult *r, int64_t *nrows);                  /* all partitions (groups after the merge) */
QE_API int qe_mresult_num_columns(const qe_mresult *r, int *ncols);
QE_API qe_result *qe_mresult_part(qe_mresult *r, int part);
QE_API int qe_mresult_part_rows(qe_mresult *r, int part, int64_t *nrows);
/* partitions concatenated in order; values / validity as qe_result_column */
QE_API int qe_mresult_column(qe_mresult *r, int k, void *dst_values, uint8_t *dst_validity);
/* global row indices (the partition's first row added), 64-bit */
QE_API int qe_mresult_selection(qe_mresult *r, uint64_t *dst);
QE_API int qe_mresult_agg(qe_mresult *r, int k, double *value, int *is_null);
QE_API int qe_mresult_destroy(qe_mresult *r);

/* Diagnostics, no GPU needed: bind `programs` (filter first when has_filter,
 * then nproj projections) against a schema, write the generated CUDA source
 * into src_buf and, when compile != 0, compile it with NVRTC for sm_100a
 * (cubin copied into cubin_buf when it fits; its size into *cubin_bytes).
 * steps / vec: tile shape overrides, 0 = the library's choice; staging: as the "staging" option. */
QE_API int qe_debug_generate(const qe_col_desc *cols, int ncols,
                             const uint32_t *const *programs, const size_t *nwords,
                             int has_filter, int nproj, const int32_t *agg_fns, int nagg,
                             int steps, int vec, int staging, int compile,
                             char *src_buf, size_t src_cap, char *cubin_buf, size_t cubin_cap, size_t *cubin_bytes);

/* Same for a group-by: programs = [filter,] nkeys key programs, nagg aggregate inputs; atomic: 0 private, 1 global atomics, 2 CTA-shared (one copy), 3 registers, 4 lane copies (32), 5 eight copies, 6 one copy per warp, 7 / 8 / 9 sixteen / four / two copies, 10 / 11 the two passes (scatter, gather). */
QE_API int qe_debug_generate_groupby(const qe_col_desc *cols, int ncols,
                                     const uint32_t *const *programs, const size_t *nwords,
                                     int has_filter, int nkeys, const int32_t *agg_fns, int nagg, int atomic, int compile,
                                     char *src_buf, size_t src_cap, size_t *cubin_bytes);

/* Ahead-of-time compilation, no GPU needed: compile the kernel a run of this pipeline would ask for (context options at
 * their defaults) through the cubin caches, so that it lands in $QE_CACHE_DIR; queryengine_b200/build.py bakes the
 * files into the library.  programs = [filter,] projections -- or, with group_keys >= 0, [filter,] keys then aggregate
 * inputs (nproj = keys + inputs; the register variant of the group-by).  key_out receives the 32-digit cache key. */
typedef struct qe_aot_desc {
    const qe_col_desc *cols;
    int32_t ncols;
    const uint32_t *const *programs;
    const size_t *nwords;
    int32_t has_filter, nproj;
    const int32_t *agg_fns;
    int32_t nagg;
    int32_t group_keys;                /* -1: not a group-by */
    double sel_hint;                   /* filter: selectivity the same query measured on its previous run, or -1 */
    int32_t min_blocks;                /* dense projection: 4 = the variant tried first, 0 = the fallback */
    int32_t seen_groups;               /* group-by: groups the same query found on its previous run, or 0 */
    uint32_t run_flags;                /* QE_RUN_NO_VALUES */
} qe_aot_desc;
QE_API int qe_aot_compile(const qe_aot_desc *desc, char key_out[33], int *origin);

/* ------------------------------------------------------------------ no device allocation
 * NO COUNTERPART IN THE REFERENCE: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 *
 * SUM / COUNT / MIN / MAX over the non-null rows of one DOUBLE column that is already on the device: the global
 * aggregation's fixed shape (sub-tiles of 16 384 rows, 64 of them per super-tile of QE_AGG_SUPER_ROWS rows, super-tiles
 * in row order; -fmad=false), so the result does not depend on the grid.  The per-sub-tile and per-super-tile partials --
 * the scratch memory qe_pipeline_run takes from the context's pool -- live in one file-scope __device__ array of
 * QE_NOALLOC_SCRATCH_BYTES instead: every GPU that loads the library holds that much until the process exits.  The call
 * allocates nothing (no cudaMalloc / cudaFree, no context, no pool) and only enqueues three kernels on `stream`; it does not
 * synchronize.
 *   device      CUDA device index that owns d_values / d_validity / d_out (made current for the call, then restored)
 *   d_values    nrows doubles, 16-byte aligned              d_validity  Arrow bitmap (bit i = row i non-null) or NULL
 *   d_out       4 doubles, written by the last kernel: [0] SUM (0.0 without rows), [1] COUNT of non-null rows as a double
 *               (exact below 2^53), [2] MIN (+Infinity without rows), [3] MAX (-Infinity without rows); Math.min / Math.max
 *               and unfused adds as everywhere else in the library
 *   stream      cudaStream_t, or NULL for the default stream
 * nrows > QE_NOALLOC_MAX_ROWS is refused with QE_ERR_INVALID_ARGUMENT (the array has a fixed size; split the column or use
 * qe_pipeline_run).  EVERY call on a device shares the array and the array keeps the last call's contents: a call writes
 * every word before it reads it, so nothing needs clearing, but two calls that overlap on the GPU (different streams, or two
 * host threads) RACE for it -- keep calls to this function on one stream per device, or order them with events. */
#define QE_NOALLOC_MAX_ROWS 268435456LL        /* 2^28 rows = 2 GiB of doubles */
#define QE_NOALLOC_SCRATCH_BYTES 532480        /* (16 384 sub-tile + 256 super-tile partials) x 32 B */
QE_API int qe_agg_column_noalloc(int device, const double *d_values, const uint8_t *d_validity, int64_t nrows,
                                 double *d_out, void *stream);

QE_API const char *qe_last_error(void);
QE_API const char *qe_version(void);

#ifdef __cplusplus
}
#endif
#endif /* QEGPU_H */
