// qe_device.cuh -- hand-written sm_100a kernel skeletons of the filter/scan hot path.
//
// Three kernel bodies share one structure (the __global__ entry points that call
// them are in the generated source and in qe_kernels.cu).  Each is a template
// over a "pipeline program" P that supplies the per-row work (load the referenced columns,
// evaluate the flattened expression trees in registers, say what to emit):
//
//   qe_filter_body<P>     Scan -> Filter -> [Projection]  -> selection vector +
//                         order-preserving compacted outputs      (K1)
//   qe_project_body<P>    Scan -> Projection (no filter)  -> dense outputs +
//                         validity words                           (K2)
//   qe_agg_body<P>        Scan -> [Filter] -> Projection -> GlobalAggregation
//                         -> per-CTA partials, fixed-order final reduce (K3)
//
// P is either generated per query (qe_codegen.cpp emits straight-line code and
// NVRTC compiles it together with this header -- the B200 counterpart of the
// reference's JVM bytecode back end, M/evaluator/BytecodeCompiler.kt:134-175)
// or the generic interpreter in qe_interp.cuh.
//
// Replaces, per row: RowCallable.invoke (M/evaluator/Compiler.kt:9-11),
// FilterOperator.next (M/operator/FilterOperator.kt:14-25), ProjectionOperator.next
// (M/operator/ProjectionOperator.kt:15-19), MemorySourceOperator.next
// (M/operator/MemorySourceOperator.kt:18-36) and GlobalAggregationOperator.open
// (M/operator/GlobalAggregationOperator.kt:10-25).
//
// Row -> thread mapping.  A CTA has QE_BLOCK = 256 threads (8 warps).  A tile is
// 8 warps x STEPS steps x (32 lanes x VEC rows).  Warp w owns the contiguous
// rows [w*STEPS*32*VEC, (w+1)*STEPS*32*VEC) of the tile; in step s lane l holds
// the VEC adjacent rows starting at  s*32*VEC + l*VEC, so one step of one warp
// reads 32*VEC*8 contiguous bytes per 8-byte column with one 128-bit (VEC=2) or
// 256-bit (VEC=4) load per lane, and VEC*32 validity bits = whole 32-bit words.
//
// This header must compile under NVRTC: no standard headers, no host code.
#ifndef QE_DEVICE_CUH
#define QE_DEVICE_CUH

#include "qe_ops.h"

#ifndef QE_DEBUG_FLAGS
#define QE_DEBUG_FLAGS 0      // timing experiments only (see qe_filter_body and the ring body; two-pass group-by: 256 pass 2 without
                              // the accumulator updates, 512 pass 2 loads only, 1024 pass 1 without appends, 2048 pass 1 without flushes)
#endif

#define QE_BLOCK 256
#define QE_WARPS 8
#define QE_FULL_MASK 0xffffffffu

#define QE_MAX_COLS 16      // referenced scan columns per pipeline
#define QE_MAX_OUTS 8       // projection outputs per pipeline
#define QE_MAX_AGGS 8
#define QE_MAX_LITS 24      // literal parameter slots per pipeline (qe_program.h)

// Global aggregation is folded in a fixed shape over the ROW INDEX, not over the grid: every CTA owns whole
// sub-tiles of QE_AGG_SUB_ROWS rows (one partial each), QE_AGG_SUPER_ROWS / QE_AGG_SUB_ROWS consecutive sub-tile
// partials are folded by one warp in a fixed tree (qe_agg_final_kernel), and the super-tile partials are added up in
// ascending row order on the host.  SUM / AVG therefore do not depend on the grid size, on the number of streamed
// chunks or on the number of GPUs, as long as every partition / chunk starts at a multiple of QE_AGG_SUPER_ROWS.
#define QE_AGG_SUB_ROWS 16384
#define QE_AGG_SUPER_ROWS 1048576

// physical column types (values match include/qegpu.h)
#define QE_T_STR 0      /* dictionary codes, int32 */
#define QE_T_F64 1
#define QE_T_BOOL 2
#define QE_T_I64 3
#define QE_T_DATE32 4

// aggregation functions (ordinals of the reference's AggregationFunction enum,
// M/ast/Functions.kt:24-26)
#define QE_AGG_MIN 0
#define QE_AGG_MAX 1
#define QE_AGG_SUM 2
#define QE_AGG_COUNT 3
#define QE_AGG_AVG 4

// Kernel arguments.  Column pointers are already offset to the first row of the
// batch; every column / bitmap allocation is zero-padded to a whole number of
// tiles, so loads never need a bounds check -- only the results are masked with
// `row < nrows`.
struct QeParams {
    const void   *col_values[QE_MAX_COLS];
    const qe_u32 *col_valid[QE_MAX_COLS];     // Arrow validity bitmap (LSB first, 1 = non-null), 0 = non-nullable
    void         *out_values[QE_MAX_OUTS];    // f64 / i64 / i32 arrays, or bitmap words for BOOLEAN outputs
    qe_u32       *out_valid[QE_MAX_OUTS];     // validity bitmap words of the output, 0 = statically non-null
    qe_u32       *sel;                        // selection vector (row indices relative to the table), may be 0
    qe_u64       *tile_state;                 // decoupled-lookback descriptors, one per tile, zeroed per run
    qe_u32       *ticket;                     // dynamic tile counter, zeroed per run
    qe_u64       *out_count;                  // number of rows kept
    const qe_u64 *out_base;                   // append mode: rows already in the outputs (device word), 0 = none
    qe_u64        out_base_mask;              // bits of *out_base that offset this run's outputs: all ones = append to one
                                              // buffer; 7 = a streamed chunk with buffers of its own keeps the bit phase only
    double       *agg_partials;               // [sub-tiles][aggregates][4] = sum, count (bits), min, max
    const qe_u32 *iprog;                      // interpreter only: serialized pipeline (qe_interp.cuh)
    qe_u64       *debug;                      // [4096][8] per-CTA phase cycle counters (QE_DEBUG_FLAGS & 16)
    qe_i64        nrows;                      // rows in this batch
    qe_i64        out_capacity;               // rows the compacted outputs can hold
    qe_u32        ntiles;
    qe_u32        sel_base;                   // table row index of the batch's first row
    // group-by (qe_groupby_*_body): open-addressing table of the groups seen so far, in HBM
    qe_u32       *g_state;                    // [slots] 0 empty, 1 key being written, 2 ready
    qe_u64       *g_keys;                     // [slots][NKEYS] canonical key bits
    qe_u32       *g_nulls;                    // [slots] bit i: key i is null
    qe_u64       *g_acc;                      // [slots][NWORDS] accumulator words (word 0 = first row << 32 | rows)
    qe_u32       *g_ctl;                      // [0] groups claimed  [1] overflow flags (1 private capacity, 2 table capacity)
    qe_u64       *g_part;                     // private variant: [grid][LMAX][NWORDS] per-CTA partial accumulators
    qe_u32       *g_part_slot;                // private variant: [grid][1 + LMAX]: local groups, then the table slot of each
    qe_u32        g_slot_mask;                // slots - 1 (a power of two)
    qe_u32        g_max_groups;
    // group-by in two passes (qe_groupby_scatter_body / qe_groupby_gather_body): records partitioned by key hash
    qe_u64       *pt_words;                   // [RW][NPART * pt_cap]: word w of every record, partition q at [q * pt_cap, (q + 1) * pt_cap)
    qe_u32       *pt_cursor;                  // [NPART] records written to every partition
    qe_u32        pt_cap;                     // records a partition holds
    qe_u32        pt_pad;
    // literals of the expression trees (generated code reads the values from here: one kernel per literal CLASS)
    double        lit_f64[QE_MAX_LITS];       // NumericLiteralExpression values
    qe_i64        lit_lo[QE_MAX_LITS];        // floor(k) for integer-domain comparisons; dictionary code of a string literal
    qe_i64        lit_hi[QE_MAX_LITS];        // ceil(k)
    double        lit_abs[QE_MAX_LITS];       // |k|      (MOD by a literal, qe_mod_k)
    double        lit_inv[QE_MAX_LITS];       // 1 / |k|
};

// rows already in the outputs before this run (append mode), and the part of it that offsets this run's positions
__device__ __forceinline__ qe_u64 qe_out_base_full(const QeParams &p) { return p.out_base ? *p.out_base : 0ull; }


// ---------------------------------------------------------------- memory ops

// Streaming loads: read-only path, no L1 allocation (every byte is used once).
__device__ __forceinline__ void qe_ldg_f64x2(const double *p, double &a, double &b)
{
    asm volatile("ld.global.nc.L1::no_allocate.v2.f64 {%0,%1}, [%2];" : "=d"(a), "=d"(b) : "l"(p));
}

// 256-bit load (new in sm_100: LDG.E.256), with an evict-first L2 hint
__device__ __forceinline__ void qe_ldg_f64x4(const double *p, double &a, double &b, double &c, double &d)
{
    asm volatile("ld.global.nc.L1::no_allocate.L2::evict_first.v4.f64 {%0,%1,%2,%3}, [%4];"
                 : "=d"(a), "=d"(b), "=d"(c), "=d"(d) : "l"(p));
}

__device__ __forceinline__ void qe_ldg_i64x2(const qe_i64 *p, qe_i64 &a, qe_i64 &b)
{
    asm volatile("ld.global.nc.L1::no_allocate.v2.s64 {%0,%1}, [%2];" : "=l"(a), "=l"(b) : "l"(p));
}

__device__ __forceinline__ void qe_ldg_i64x4(const qe_i64 *p, qe_i64 &a, qe_i64 &b, qe_i64 &c, qe_i64 &d)
{
    asm volatile("ld.global.nc.L1::no_allocate.L2::evict_first.v4.s64 {%0,%1,%2,%3}, [%4];"
                 : "=l"(a), "=l"(b), "=l"(c), "=l"(d) : "l"(p));
}

__device__ __forceinline__ void qe_ldg_i32x2(const qe_i32 *p, qe_i32 &a, qe_i32 &b)
{
    asm volatile("ld.global.nc.L1::no_allocate.v2.s32 {%0,%1}, [%2];" : "=r"(a), "=r"(b) : "l"(p));
}

__device__ __forceinline__ void qe_ldg_i32x4(const qe_i32 *p, qe_i32 &a, qe_i32 &b, qe_i32 &c, qe_i32 &d)
{
    asm volatile("ld.global.nc.L1::no_allocate.v4.s32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(a), "=r"(b), "=r"(c), "=r"(d) : "l"(p));
}

__device__ __forceinline__ qe_u32 qe_ldg_u32(const qe_u32 *p)
{
    qe_u32 v;
    asm volatile("ld.global.nc.u32 %0, [%1];" : "=r"(v) : "l"(p));
    return v;
}

// Lookback descriptors: value and status share one 64-bit word, so relaxed
// gpu-scope accesses are enough (nothing else is published through them).
__device__ __forceinline__ qe_u64 qe_ld_relaxed(const qe_u64 *p)
{
    qe_u64 v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}

__device__ __forceinline__ void qe_st_relaxed(qe_u64 *p, qe_u64 v)
{
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" :: "l"(p), "l"(v) : "memory");
}

// ------------------------------------------------------- column load helpers
// `r` is the first of the VEC adjacent rows this lane holds (a multiple of VEC).

template <int VEC>
__device__ __forceinline__ void qe_load_f64(const void *col, qe_i64 r, double (&v)[VEC])
{
    const double *p = (const double *)col + r;
    if constexpr (VEC == 4) qe_ldg_f64x4(p, v[0], v[1], v[2], v[3]);
    else qe_ldg_f64x2(p, v[0], v[1]);
}

template <int VEC>
__device__ __forceinline__ void qe_load_i64(const void *col, qe_i64 r, qe_i64 (&v)[VEC])
{
    const qe_i64 *p = (const qe_i64 *)col + r;
    if constexpr (VEC == 4) qe_ldg_i64x4(p, v[0], v[1], v[2], v[3]);
    else qe_ldg_i64x2(p, v[0], v[1]);
}

template <int VEC>
__device__ __forceinline__ void qe_load_i32(const void *col, qe_i64 r, qe_i32 (&v)[VEC])
{
    const qe_i32 *p = (const qe_i32 *)col + r;
    if constexpr (VEC == 4) qe_ldg_i32x4(p, v[0], v[1], v[2], v[3]);
    else qe_ldg_i32x2(p, v[0], v[1]);
}

// The VEC bits of rows r .. r+VEC-1 of a bitmap.  The lanes of a warp ask for
// VEC*32 consecutive bits = VEC/… whole 32-bit words; lanes that share a word
// issue the same address, which the LSU broadcasts.
template <int VEC>
__device__ __forceinline__ qe_u32 qe_load_bits(const qe_u32 *bitmap, qe_i64 r)
{
    qe_u32 w = qe_ldg_u32(bitmap + (r >> 5));
    return (w >> ((qe_u32)r & 31u)) & ((1u << VEC) - 1u);
}

// Dense output stores: the VEC adjacent values of one lane, one 128/256-bit store.
template <int VEC>
__device__ __forceinline__ void qe_store_f64(void *col, qe_i64 r, const double (&v)[VEC])
{
    double *p = (double *)col + r;
    if constexpr (VEC == 4)
        asm volatile("st.global.v4.f64 [%0], {%1,%2,%3,%4};" :: "l"(p), "d"(v[0]), "d"(v[1]), "d"(v[2]), "d"(v[3]) : "memory");
    else
        asm volatile("st.global.v2.f64 [%0], {%1,%2};" :: "l"(p), "d"(v[0]), "d"(v[1]) : "memory");
}

template <int VEC>
__device__ __forceinline__ void qe_store_i64(void *col, qe_i64 r, const qe_i64 (&v)[VEC])
{
    qe_i64 *p = (qe_i64 *)col + r;
    if constexpr (VEC == 4)
        asm volatile("st.global.v4.s64 [%0], {%1,%2,%3,%4};" :: "l"(p), "l"(v[0]), "l"(v[1]), "l"(v[2]), "l"(v[3]) : "memory");
    else
        asm volatile("st.global.v2.s64 [%0], {%1,%2};" :: "l"(p), "l"(v[0]), "l"(v[1]) : "memory");
}

template <int VEC>
__device__ __forceinline__ void qe_store_i32(void *col, qe_i64 r, const qe_i32 (&v)[VEC])
{
    qe_i32 *p = (qe_i32 *)col + r;
    if constexpr (VEC == 4)
        asm volatile("st.global.v4.s32 [%0], {%1,%2,%3,%4};" :: "l"(p), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]) : "memory");
    else
        asm volatile("st.global.v2.s32 [%0], {%1,%2};" :: "l"(p), "r"(v[0]), "r"(v[1]) : "memory");
}

// ------------------------------------------------------------ warp utilities

__device__ __forceinline__ qe_u32 qe_lane() { return threadIdx.x & 31u; }
__device__ __forceinline__ qe_u32 qe_warp() { return threadIdx.x >> 5; }

__device__ __forceinline__ qe_u32 qe_lanemask_lt()
{
    qe_u32 m;
    asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
    return m;
}

__device__ __forceinline__ double qe_shfl_down_f64(double v, int d)
{
    return __shfl_down_sync(QE_FULL_MASK, v, d);
}

// Validity / boolean words of one warp step (dense output).  Lane l holds VEC
// bits for rows l*VEC ..; the warp assembles the VEC 32-bit words with one
// REDUX.OR each and lane 0 stores them.
template <int VEC>
__device__ __forceinline__ void qe_store_bits_dense(qe_u32 *bitmap, qe_i64 warp_row0, qe_u32 bits)
{
    const qe_u32 q = qe_lane() * VEC;                 // bit position inside the warp step
    qe_u32 words[VEC];
#pragma unroll
    for (int h = 0; h < VEC; h++)
        words[h] = __reduce_or_sync(QE_FULL_MASK, (q >> 5) == (qe_u32)h ? bits << (q & 31u) : 0u);
    if (qe_lane() == 0) {
        qe_u32 *dst = bitmap + (warp_row0 >> 5);
        if constexpr (VEC == 4) {
            asm volatile("st.global.v4.u32 [%0], {%1,%2,%3,%4};"
                         :: "l"(dst), "r"(words[0]), "r"(words[1]), "r"(words[2]), "r"(words[3]) : "memory");
        } else {
            asm volatile("st.global.v2.u32 [%0], {%1,%2};" :: "l"(dst), "r"(words[0]), "r"(words[1]) : "memory");
        }
    }
}

// Compacted bit column: the kept rows of one warp step occupy output bits
// [gpos, gpos + n) with n <= 32*VEC.  Every kept lane drops its bit at its rank,
// the warp ORs the words together and lanes 0..VEC OR them into the (pre-zeroed)
// output bitmap.  Tiles meet inside a word, hence the atomics; OR makes the
// result independent of arrival order.
template <int VEC>
__device__ __forceinline__ void qe_store_bits_compact(qe_u32 *bitmap, qe_u64 gpos, const qe_u32 (&rank)[VEC],
                                                      qe_u32 keep, qe_u32 bits)
{
    qe_u32 words[VEC + 1];
#pragma unroll
    for (int h = 0; h < VEC; h++) {
        qe_u32 mine = 0;
#pragma unroll
        for (int j = 0; j < VEC; j++) {
            const bool on = ((keep >> j) & 1u) && ((bits >> j) & 1u) && (rank[j] >> 5) == (qe_u32)h;
            mine |= on ? 1u << (rank[j] & 31u) : 0u;
        }
        words[h] = __reduce_or_sync(QE_FULL_MASK, mine);
    }
    words[VEC] = 0;
    const qe_u32 sh = (qe_u32)gpos & 31u;
    const qe_u32 l = qe_lane();
    if (l <= VEC) {
        // word l of the shifted bit string = (words[l] << sh) | (words[l-1] >> (32 - sh))
        qe_u32 cur = 0, prev = 0;
#pragma unroll
        for (int h = 0; h <= VEC; h++) {
            if ((qe_u32)h == l) cur = words[h];
            if ((qe_u32)h + 1 == l) prev = words[h];
        }
        const qe_u32 v = (cur << sh) | (sh ? prev >> (32u - sh) : 0u);
        if (v) atomicOr(bitmap + (gpos >> 5) + l, v);
    }
}

// --------------------------------------------------------- decoupled lookback
// Single-pass prefix over the per-tile kept-row counts (Merrill & Garland).
// Descriptor = status (2 bits) | value (62 bits).  Tiles are handed out through
// an atomic ticket, so every predecessor of a tile is already resident or done
// and the spin below cannot deadlock.

#define QE_TS_INVALID   0ull
#define QE_TS_AGGREGATE 1ull
#define QE_TS_PREFIX    2ull
#define QE_TS_VALUE_MASK ((1ull << 62) - 1ull)

// Called by warp 0.  Returns the number of kept rows in all earlier tiles.
__device__ __forceinline__ qe_u64 qe_lookback(qe_u64 *state, qe_u32 tile, qe_u32 tile_total)
{
    const qe_u32 lane = qe_lane();
    if (tile == 0) {
        if (lane == 0) qe_st_relaxed(state, (QE_TS_PREFIX << 62) | (qe_u64)tile_total);
        return 0;
    }
    if (lane == 0) qe_st_relaxed(state + tile, (QE_TS_AGGREGATE << 62) | (qe_u64)tile_total);

    qe_u64 exclusive = 0;
    qe_i64 look = (qe_i64)tile - 1;
    for (;;) {
        const qe_i64 idx = look - (qe_i64)lane;
        qe_u64 d = QE_TS_PREFIX << 62;                       // before tile 0: prefix 0
        if (idx >= 0) {
            do { d = qe_ld_relaxed(state + idx); } while ((d >> 62) == QE_TS_INVALID);
        }
        const qe_u32 has_prefix = __ballot_sync(QE_FULL_MASK, (d >> 62) == QE_TS_PREFIX);
        const qe_u32 first = has_prefix ? (qe_u32)__ffs((int)has_prefix) - 1u : 32u;
        qe_u64 v = lane <= first ? (d & QE_TS_VALUE_MASK) : 0ull;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(QE_FULL_MASK, v, o);
        exclusive += v;
        if (has_prefix) break;
        look -= 32;
    }
    if (lane == 0) qe_st_relaxed(state + tile, (QE_TS_PREFIX << 62) | (exclusive + (qe_u64)tile_total));
    return exclusive;
}

// Block-wide look-back.  A persistent grid keeps every resident CTA in the same
// phase, so when a tile starts looking back none of the ~600 tiles in flight has
// its prefix yet and a warp-wide window (32 descriptors per L2 round trip) walks
// ~20 windows = 7 us per tile (measured).  Here all 256 threads look at
// 256 * QE_LB_PER_THREAD predecessors per round trip; warp w covers the w-th
// nearest run of 32 * QE_LB_PER_THREAD tiles.  Called by every thread of the CTA
// (contains block barriers); returns the exclusive prefix to all of them.
#define QE_LB_PER_THREAD 2

// one thread: make the tile's kept-row count visible to its successors
__device__ __forceinline__ void qe_publish_aggregate(qe_u64 *state, qe_u32 tile, qe_u32 tile_total)
{
    qe_st_relaxed(state + tile, ((tile == 0 ? QE_TS_PREFIX : QE_TS_AGGREGATE) << 62) | (qe_u64)tile_total);
}

template <bool PUBLISH_FIRST = true>
__device__ __forceinline__ qe_u64 qe_block_lookback(qe_u64 *state, qe_u32 tile, qe_u32 tile_total,
                                                    qe_u64 *s_sum, qe_u32 *s_found)
{
    const qe_u32 lane = qe_lane(), warp = qe_warp();
    if (PUBLISH_FIRST && threadIdx.x == 0) qe_publish_aggregate(state, tile, tile_total);
    if (tile == 0) {
        __syncthreads();                                      // callers rely on one barrier in here
        return 0;
    }

    qe_u64 exclusive = 0;
    qe_i64 look = (qe_i64)tile - 1;
    for (;;) {
        qe_u64 d[QE_LB_PER_THREAD];
        qe_i64 idx[QE_LB_PER_THREAD];
#pragma unroll
        for (int k = 0; k < QE_LB_PER_THREAD; k++) {          // all reads in flight together
            idx[k] = look - (qe_i64)(warp * 32 * QE_LB_PER_THREAD + k * 32 + lane);
            d[k] = idx[k] >= 0 ? qe_ld_relaxed(state + idx[k]) : (QE_TS_PREFIX << 62);
        }
        qe_u64 wsum = 0;
        bool wfound = false;
#pragma unroll
        for (int k = 0; k < QE_LB_PER_THREAD; k++) {
            if (!wfound) {                                    // warp-uniform
                while ((d[k] >> 62) == QE_TS_INVALID) d[k] = qe_ld_relaxed(state + idx[k]);
                const qe_u32 has_prefix = __ballot_sync(QE_FULL_MASK, (d[k] >> 62) == QE_TS_PREFIX);
                const qe_u32 first = has_prefix ? (qe_u32)__ffs((int)has_prefix) - 1u : 32u;
                qe_u64 v = lane <= first ? (d[k] & QE_TS_VALUE_MASK) : 0ull;
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(QE_FULL_MASK, v, o);
                wsum += v;
                wfound = has_prefix != 0;
            }
        }
        if (lane == 0) {
            s_sum[warp] = wsum;
            s_found[warp] = wfound ? 1u : 0u;
        }
        __syncthreads();
        bool found = false;
#pragma unroll
        for (int w = 0; w < QE_WARPS; w++) {
            if (!found) {
                exclusive += s_sum[w];
                found = s_found[w] != 0;
            }
        }
        if (found) break;
        look -= QE_WARPS * 32 * QE_LB_PER_THREAD;
        __syncthreads();                                      // s_sum / s_found are rewritten next round
    }
    if (threadIdx.x == 0) qe_st_relaxed(state + tile, (QE_TS_PREFIX << 62) | (exclusive + (qe_u64)tile_total));
    return exclusive;
}

// ------------------------------------------------------------- accumulators
// One per aggregate, per thread.  Only the fields the function needs are
// touched (FN is a compile-time constant), so the rest folds away.
struct QeAcc {
    double sum;
    qe_i64 count;
    double mn;
    double mx;
};

__device__ __forceinline__ void qe_acc_init(QeAcc &a)
{
    a.sum = 0.0;
    a.count = 0;
    a.mn = qe_bits_f64((qe_i64)0x7ff0000000000000LL);       // +Infinity, M/evaluator/Accumulators.kt:57
    a.mx = qe_bits_f64((qe_i64)0xfff0000000000000ULL);      // -Infinity, :75
}

// accumulate(value) when the value is non-null and the row passed the filter
// (M/operator/GlobalAggregationOperator.kt:17-20)
template <int FN>
__device__ __forceinline__ void qe_acc_add(QeAcc &a, double v, bool on)
{
    if (on) {
        a.count += 1;
        if (FN == QE_AGG_SUM || FN == QE_AGG_AVG) a.sum += v;
        if (FN == QE_AGG_MIN) a.mn = qe_math_min(a.mn, v);
        if (FN == QE_AGG_MAX) a.mx = qe_math_max(a.mx, v);
    }
}

__device__ __forceinline__ qe_i64 qe_shfl_down_i64(qe_i64 v, int d)
{
    return __shfl_down_sync(QE_FULL_MASK, v, d);
}

// Merge `b` into `a` (fixed operand order: a is the lower-numbered lane / warp / CTA)
template <int FN>
__device__ __forceinline__ void qe_acc_merge(QeAcc &a, const QeAcc &b)
{
    a.count += b.count;
    if (FN == QE_AGG_SUM || FN == QE_AGG_AVG) a.sum += b.sum;
    if (FN == QE_AGG_MIN) a.mn = qe_math_min(a.mn, b.mn);
    if (FN == QE_AGG_MAX) a.mx = qe_math_max(a.mx, b.mx);
}

template <int FN>
__device__ __forceinline__ void qe_acc_warp_reduce(QeAcc &a)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        QeAcc b;
        b.count = qe_shfl_down_i64(a.count, o);
        b.sum = (FN == QE_AGG_SUM || FN == QE_AGG_AVG) ? qe_shfl_down_f64(a.sum, o) : 0.0;
        b.mn = (FN == QE_AGG_MIN) ? qe_shfl_down_f64(a.mn, o) : a.mn;
        b.mx = (FN == QE_AGG_MAX) ? qe_shfl_down_f64(a.mx, o) : a.mx;
        qe_acc_merge<FN>(a, b);
    }
}

// Block reduce in fixed order (lane tree, then warps 0..7 sequentially) and
// write this CTA's partial.  `slot` is a [QE_WARPS] shared array.
template <int FN>
__device__ __forceinline__ void qe_acc_block_reduce_store(QeAcc a, QeAcc *slot, double *partials, qe_i64 part, int nagg, int k)
{
    qe_acc_warp_reduce<FN>(a);
    __syncthreads();
    if (qe_lane() == 0) slot[qe_warp()] = a;
    __syncthreads();
    if (threadIdx.x == 0) {
        QeAcc t = slot[0];
        for (int w = 1; w < QE_WARPS; w++) qe_acc_merge<FN>(t, slot[w]);
        double *dst = partials + (part * nagg + k) * 4;
        dst[0] = t.sum;
        dst[1] = qe_bits_f64(t.count);
        dst[2] = t.mn;
        dst[3] = t.mx;
    }
}

// Field-agnostic merge (every field folded; fields an aggregate does not use
// hold their init values or are ignored by the host).  Used by the interpreter
// and by the final reduce of the per-CTA partials.
__device__ __forceinline__ void qe_acc_merge_all(QeAcc &a, const QeAcc &b)
{
    a.count += b.count;
    a.sum += b.sum;
    a.mn = qe_math_min(a.mn, b.mn);
    a.mx = qe_math_max(a.mx, b.mx);
}

// Same block reduce as qe_acc_block_reduce_store, all four fields.
__device__ __forceinline__ void qe_acc_block_reduce_store_all(QeAcc a, QeAcc *slot, double *partials, qe_i64 part, int nagg, int k)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        QeAcc b;
        b.count = qe_shfl_down_i64(a.count, o);
        b.sum = qe_shfl_down_f64(a.sum, o);
        b.mn = qe_shfl_down_f64(a.mn, o);
        b.mx = qe_shfl_down_f64(a.mx, o);
        qe_acc_merge_all(a, b);
    }
    __syncthreads();
    if (qe_lane() == 0) slot[qe_warp()] = a;
    __syncthreads();
    if (threadIdx.x == 0) {
        QeAcc t = slot[0];
        for (int w = 1; w < QE_WARPS; w++) qe_acc_merge_all(t, slot[w]);
        double *dst = partials + (part * nagg + k) * 4;
        dst[0] = t.sum;
        dst[1] = qe_bits_f64(t.count);
        dst[2] = t.mn;
        dst[3] = t.mx;
    }
}

// ------------------------------------------------------------------ kernels
//
// What a pipeline program P provides:
//   static constexpr int STEPS, VEC, NBITCOLS;
//   struct Row;   inputs of one row, in registers
//   struct Out;   outputs of one row (projection values), in registers
//   static void load(const QeParams&, qe_i64 r, Row (&rows)[VEC]);      r = lane's first row
//   static bool eval(const QeParams&, const Row&, Out&);               returns the filter verdict
//                                                                      (non-null AND true, FilterOperator.kt:20)
//   filter kernel:
//     static void emit(const QeParams&, qe_u64 pos, const Out&);       value outputs at compacted position
//     static qe_u32* bitcol(const QeParams&, int b);                   b-th compacted bit column (or 0)
//     static bool bit(const Out&, int b);
//   project kernel:
//     static void emit_dense(const QeParams&, qe_i64 r, const Out (&)[VEC], qe_u32 inb);
//     bitcol / bit as above (dense words)
//   agg kernel:
//     struct Accs; static void acc_init(Accs&); static void acc_add(const QeParams&, Accs&, const Out&, bool keep);
//     static void acc_store(const QeParams&, Accs&, QeAcc *slot);

template <class P>
__device__ __forceinline__ void qe_tile_eval(const QeParams &p, qe_i64 warp_row0,
                                             typename P::Out (&outs)[P::STEPS][P::VEC],
                                             qe_u32 (&keep)[P::STEPS], qe_u32 (&inb)[P::STEPS])
{
    constexpr int STEPS = P::STEPS, VEC = P::VEC;
    typename P::Row rows[STEPS][VEC];
    const qe_u32 lane = qe_lane();
    // issue every load of the tile first: STEPS x (columns) independent 16/32-byte
    // requests per thread in flight before the first use
#pragma unroll
    for (int s = 0; s < STEPS; s++)
        P::load(p, warp_row0 + (qe_i64)s * 32 * VEC + lane * VEC, rows[s]);
#pragma unroll
    for (int s = 0; s < STEPS; s++) {
        const qe_i64 r = warp_row0 + (qe_i64)s * 32 * VEC + lane * VEC;
        keep[s] = 0;
        inb[s] = 0;
#pragma unroll
        for (int j = 0; j < VEC; j++) {
            const bool in = r + j < p.nrows;
            const bool k = P::eval(p, rows[s][j], outs[s][j]) && in;
            keep[s] |= (k ? 1u : 0u) << j;
            inb[s] |= (in ? 1u : 0u) << j;
        }
    }
}

// Evaluate rows that are already in registers (used by both filter bodies).
template <class P>
__device__ __forceinline__ void qe_rows_eval(const QeParams &p, qe_i64 warp_row0,
                                             const typename P::Row (&rows)[P::STEPS][P::VEC],
                                             typename P::Out (&outs)[P::STEPS][P::VEC],
                                             qe_u32 (&keep)[P::STEPS])
{
    constexpr int STEPS = P::STEPS, VEC = P::VEC;
    const qe_u32 lane = qe_lane();
#pragma unroll
    for (int s = 0; s < STEPS; s++) {
        const qe_i64 r = warp_row0 + (qe_i64)s * 32 * VEC + lane * VEC;
        keep[s] = 0;
#pragma unroll
        for (int j = 0; j < VEC; j++) {
            const bool k = P::eval(p, rows[s][j], outs[s][j]) && (r + j < p.nrows);
            keep[s] |= (k ? 1u : 0u) << j;
        }
    }
}

// Rank of every kept row inside its warp's slice (row order = step, lane, j): one
// ballot + two popcounts per (step, j).  rank0[s] = rank of the lane's first kept row
// of step s within the step; step_off[s] = kept rows of the warp in earlier steps.
template <int STEPS, int VEC>
__device__ __forceinline__ qe_u32 qe_warp_ranks(const qe_u32 (&keep)[STEPS], qe_u32 (&rank0)[STEPS], qe_u32 (&step_off)[STEPS])
{
    const qe_u32 lt = qe_lanemask_lt();
    qe_u32 wcnt = 0;
#pragma unroll
    for (int s = 0; s < STEPS; s++) {
        qe_u32 before = 0, total = 0;
#pragma unroll
        for (int j = 0; j < VEC; j++) {
            const qe_u32 b = __ballot_sync(QE_FULL_MASK, (keep[s] >> j) & 1u);
            before += __popc(b & lt);
            total += __popc(b);
        }
        rank0[s] = before;
        step_off[s] = wcnt;
        wcnt += total;
    }
    return wcnt;
}

// Ordered scatter of one warp's kept rows: a warp step's survivors land on
// consecutive addresses starting at gbase + step_off[s].
template <class P>
__device__ __forceinline__ void qe_warp_scatter(const QeParams &p, qe_i64 warp_row0, qe_u64 gbase,
                                                const typename P::Out (&outs)[P::STEPS][P::VEC],
                                                const qe_u32 (&keep)[P::STEPS], const qe_u32 (&rank0)[P::STEPS],
                                                const qe_u32 (&step_off)[P::STEPS])
{
    constexpr int STEPS = P::STEPS, VEC = P::VEC;
    const qe_u32 lane = qe_lane();
#pragma unroll
    for (int s = 0; s < STEPS; s++) {
        const qe_u64 gstep = gbase + step_off[s];
        const qe_i64 r = warp_row0 + (qe_i64)s * 32 * VEC + lane * VEC;
        qe_u32 rank[VEC];
        qe_u32 rr = rank0[s];
#pragma unroll
        for (int j = 0; j < VEC; j++) {
            rank[j] = rr;
            rr += (keep[s] >> j) & 1u;
        }
#pragma unroll
        for (int j = 0; j < VEC; j++) {
            const qe_u64 pos = gstep + rank[j];
            if (((keep[s] >> j) & 1u) && pos < (qe_u64)p.out_capacity) {
                if (p.sel) p.sel[pos] = p.sel_base + (qe_u32)(r + j);
                P::emit(p, pos, outs[s][j]);
            }
        }
        if (P::NBITCOLS > 0 && gstep < (qe_u64)p.out_capacity) {
#pragma unroll
            for (int b = 0; b < P::NBITCOLS; b++) {
                qe_u32 *bm = P::bitcol(p, b);
                if (bm) {
                    qe_u32 bits = 0;
#pragma unroll
                    for (int j = 0; j < VEC; j++) bits |= (P::bit(outs[s][j], b) ? 1u : 0u) << j;
                    qe_store_bits_compact<VEC>(bm, gstep, rank, keep[s], bits);
                }
            }
        }
    }
}

// K1, register-sourced variant: the columns are loaded straight into registers
// with 128/256-bit LDGs.  One ticket, three block barriers per tile.
template <class P>
__device__ __forceinline__ void qe_filter_body(const QeParams &p)
{
    constexpr int STEPS = P::STEPS, VEC = P::VEC;
    constexpr int WARP_ROWS = STEPS * 32 * VEC;
    constexpr int TILE = QE_WARPS * WARP_ROWS;

    __shared__ qe_u32 s_tile;
    __shared__ qe_u32 s_warp_cnt[QE_WARPS];
    __shared__ qe_u64 s_lb_sum[QE_WARPS];
    __shared__ qe_u32 s_lb_found[QE_WARPS];

    const qe_u32 lane = qe_lane(), warp = qe_warp();
    const qe_u64 base0 = qe_out_base_full(p) & p.out_base_mask;      // append mode: rows of earlier batches

#ifndef QE_DEBUG_FLAGS
#define QE_DEBUG_FLAGS 0      // timing experiments only: 1 no look-back, 2 static tiles, 4 no scatter, 8 warp look-back
#endif
    // phase timing (experiments): 0 ticket wait, 1 loads+eval, 2 barrier, 3 look-back, 4 scatter, 5 tiles
    qe_u64 dbg_acc[6] = {0, 0, 0, 0, 0, 0};
    qe_u64 dbg_last = (QE_DEBUG_FLAGS & 16) ? clock64() : 0;
#define QE_TICK(i) do { if (QE_DEBUG_FLAGS & 16) { const qe_u64 now_ = clock64(); dbg_acc[i] += now_ - dbg_last; \
                         dbg_last = now_; if ((i) == 4) dbg_acc[5] += 1; } } while (0)
    for (qe_u32 it = 0;; it++) {
        if (QE_DEBUG_FLAGS & 2) {
            if (threadIdx.x == 0) s_tile = it * gridDim.x + blockIdx.x;
        } else {
            if (threadIdx.x == 0) s_tile = atomicAdd(p.ticket, 1u);
        }
        __syncthreads();
        QE_TICK(0);
        const qe_u32 tile = s_tile;
        if (tile >= p.ntiles) break;
        const qe_i64 warp_row0 = (qe_i64)tile * TILE + (qe_i64)warp * WARP_ROWS;

        typename P::Out outs[STEPS][VEC];
        qe_u32 keep[STEPS];
        {
            typename P::Row rows[STEPS][VEC];
#pragma unroll
            for (int s = 0; s < STEPS; s++)
                P::load(p, warp_row0 + (qe_i64)s * 32 * VEC + lane * VEC, rows[s]);
            qe_rows_eval<P>(p, warp_row0, rows, outs, keep);
        }
        qe_u32 rank0[STEPS], step_off[STEPS];
        const qe_u32 wcnt = qe_warp_ranks<STEPS, VEC>(keep, rank0, step_off);
        if (lane == 0) s_warp_cnt[warp] = wcnt;
        QE_TICK(1);
        __syncthreads();
        QE_TICK(2);

        qe_u32 woff = 0, tile_total = 0;
#pragma unroll
        for (int w = 0; w < QE_WARPS; w++) {
            const qe_u32 c = s_warp_cnt[w];
            woff += (qe_u32)w < warp ? c : 0u;
            tile_total += c;
        }
        qe_u64 excl;
        if (QE_DEBUG_FLAGS & 1) {
            excl = (qe_u64)tile * (TILE / 2);
            __syncthreads();
        } else if (QE_DEBUG_FLAGS & 8) {
            if (warp == 0) {
                const qe_u64 e = qe_lookback(p.tile_state, tile, tile_total);
                if (lane == 0) s_lb_sum[0] = e;
            }
            __syncthreads();
            excl = s_lb_sum[0];
        } else {
            excl = qe_block_lookback(p.tile_state, tile, tile_total, s_lb_sum, s_lb_found);
        }
        QE_TICK(3);
        excl += base0;
        if (threadIdx.x == 0 && tile == p.ntiles - 1) *p.out_count = (qe_out_base_full(p) & ~p.out_base_mask) + excl + tile_total;
        if (!(QE_DEBUG_FLAGS & 4)) qe_warp_scatter<P>(p, warp_row0, excl + woff, outs, keep, rank0, step_off);
        QE_TICK(4);
    }
    if ((QE_DEBUG_FLAGS & 16) && threadIdx.x == 0 && blockIdx.x < 4096) {
#pragma unroll
        for (int i = 0; i < 6; i++) p.debug[blockIdx.x * 8 + i] = dbg_acc[i];
    }
}

// K1, software-pipelined variant (the default).
//
// Stable compaction has a true dependency: tile t cannot be written before every
// earlier tile has been evaluated.  Measured on B200 (tools/phase_timing.py): with
// look-back right after evaluation a CTA spends 7 000 - 15 000 cycles per tile
// polling descriptors (one L2 round trip under streaming load is ~900 cycles, and
// a persistent grid keeps all resident tiles in the same phase), while loading and
// evaluating the tile takes ~2 000.  So the look-back and scatter of tile i are
// deferred by one iteration: the CTA publishes the tile's count, claims and loads
// its next tile, and only then -- with those loads in flight -- reads the
// predecessors' descriptors (published a whole tile time ago, so there is nothing to
// spin on) and scatters tile i from registers.
template <class P>
__device__ __forceinline__ void qe_filter_pipelined_body(const QeParams &p)
{
    constexpr int STEPS = P::STEPS, VEC = P::VEC;
    constexpr int WARP_ROWS = STEPS * 32 * VEC;
    constexpr int TILE = QE_WARPS * WARP_ROWS;

    __shared__ qe_u32 s_tile[2];
    __shared__ qe_u32 s_warp_cnt[QE_WARPS];
    __shared__ qe_u64 s_lb_sum[QE_WARPS];
    __shared__ qe_u32 s_lb_found[QE_WARPS];

    const qe_u32 lane = qe_lane(), warp = qe_warp();

    if (threadIdx.x == 0) s_tile[0] = atomicAdd(p.ticket, 1u);
    __syncthreads();
    qe_u32 cur = s_tile[0];

    // state of the previous tile, kept in registers until its prefix is known
    typename P::Out prev_outs[STEPS][VEC];
    qe_u32 prev_keep[STEPS], prev_rank0[STEPS], prev_step_off[STEPS];
    qe_u32 prev_tile = 0, prev_woff = 0, prev_total = 0;
    bool have_prev = false;

    for (qe_u32 it = 0;; it++) {
        const bool have_cur = cur < p.ntiles;
        const qe_i64 warp_row0 = (qe_i64)cur * TILE + (qe_i64)warp * WARP_ROWS;
        typename P::Row rows[STEPS][VEC];
        if (have_cur) {
            // claim the tile after this one (its id is read after the next barrier) and start loading this one
            if (threadIdx.x == 0) s_tile[(it + 1) & 1] = atomicAdd(p.ticket, 1u);
#pragma unroll
            for (int s = 0; s < STEPS; s++)
                P::load(p, warp_row0 + (qe_i64)s * 32 * VEC + lane * VEC, rows[s]);
        }
        if (have_prev) {
            // the previous tile: prefix (its predecessors published long ago) and ordered scatter
            const qe_u64 excl = (qe_out_base_full(p) & p.out_base_mask) +
                                qe_block_lookback<false>(p.tile_state, prev_tile, prev_total, s_lb_sum, s_lb_found);
            if (threadIdx.x == 0 && prev_tile == p.ntiles - 1) *p.out_count = (qe_out_base_full(p) & ~p.out_base_mask) + excl + prev_total;
            const qe_i64 prev_row0 = (qe_i64)prev_tile * TILE + (qe_i64)warp * WARP_ROWS;
            qe_warp_scatter<P>(p, prev_row0, excl + prev_woff, prev_outs, prev_keep, prev_rank0, prev_step_off);
        }
        if (!have_cur) break;

        qe_rows_eval<P>(p, warp_row0, rows, prev_outs, prev_keep);
        const qe_u32 wcnt = qe_warp_ranks<STEPS, VEC>(prev_keep, prev_rank0, prev_step_off);
        if (lane == 0) s_warp_cnt[warp] = wcnt;
        __syncthreads();
        qe_u32 woff = 0, tile_total = 0;
#pragma unroll
        for (int w = 0; w < QE_WARPS; w++) {
            const qe_u32 c = s_warp_cnt[w];
            woff += (qe_u32)w < warp ? c : 0u;
            tile_total += c;
        }
        if (threadIdx.x == 0) qe_publish_aggregate(p.tile_state, cur, tile_total);
        prev_tile = cur;
        prev_woff = woff;
        prev_total = tile_total;
        have_prev = true;
        cur = s_tile[(it + 1) & 1];
        // s_warp_cnt is rewritten only after the barrier inside the next look-back
    }
}

// ---- bulk-copy (TMA) staging ---------------------------------------------------
// cp.async.bulk moves a contiguous run of a column from HBM into shared memory
// without touching the register file and signals an mbarrier when the bytes have
// landed (SASS: UBLKCP + SYNCS).  With the tile's columns landing in a ring of
// shared-memory stages, the bytes in flight per SM are bounded by shared memory
// (up to 227 KB) instead of by registers, and the loads of the next tiles overlap
// the scan / look-back / scatter of the current one.

__device__ __forceinline__ qe_u32 qe_smem_addr(const void *p)
{
    return (qe_u32)__cvta_generic_to_shared(p);
}

__device__ __forceinline__ void qe_mbar_init(qe_u32 mbar, qe_u32 count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(mbar), "r"(count) : "memory");
}

__device__ __forceinline__ void qe_mbar_expect_tx(qe_u32 mbar, qe_u32 bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(mbar), "r"(bytes) : "memory");
}

__device__ __forceinline__ void qe_mbar_wait(qe_u32 mbar, qe_u32 parity)
{
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "QE_WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra QE_DONE_%=;\n\t"
        "bra QE_WAIT_%=;\n\t"
        "QE_DONE_%=:\n\t"
        "}" :: "r"(mbar), "r"(parity) : "memory");
}

// global -> shared bulk copy; size and both addresses are multiples of 16 bytes
__device__ __forceinline__ void qe_bulk_g2s(qe_u32 dst_smem, const void *src, qe_u32 bytes, qe_u32 mbar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"(dst_smem), "l"(src), "r"(bytes), "r"(mbar) : "memory");
}

__device__ __forceinline__ void qe_lds_f64x2(const void *p, double &a, double &b)
{
    asm volatile("ld.shared.v2.f64 {%0,%1}, [%2];" : "=d"(a), "=d"(b) : "r"(qe_smem_addr(p)));
}

__device__ __forceinline__ void qe_lds_i64x2(const void *p, qe_i64 &a, qe_i64 &b)
{
    asm volatile("ld.shared.v2.s64 {%0,%1}, [%2];" : "=l"(a), "=l"(b) : "r"(qe_smem_addr(p)));
}

__device__ __forceinline__ void qe_lds_i32x2(const void *p, qe_i32 &a, qe_i32 &b)
{
    asm volatile("ld.shared.v2.s32 {%0,%1}, [%2];" : "=r"(a), "=r"(b) : "r"(qe_smem_addr(p)));
}

// VEC = 2 only: lane l reads the 16 bytes at l*16 of each 512-byte warp step (conflict-free)
__device__ __forceinline__ void qe_stage_f64(const unsigned char *col, int r, double (&v)[2])
{
    qe_lds_f64x2(col + (size_t)r * 8, v[0], v[1]);
}

__device__ __forceinline__ void qe_stage_i64(const unsigned char *col, int r, qe_i64 (&v)[2])
{
    qe_lds_i64x2(col + (size_t)r * 8, v[0], v[1]);
}

__device__ __forceinline__ void qe_stage_i32(const unsigned char *col, int r, qe_i32 (&v)[2])
{
    qe_lds_i32x2(col + (size_t)r * 4, v[0], v[1]);
}

__device__ __forceinline__ qe_u32 qe_stage_bits(const unsigned char *bitmap, int r)
{
    const qe_u32 w = *(const qe_u32 *)(bitmap + (r >> 5) * 4);
    return (w >> (r & 31)) & 3u;
}

// one row of a staged column (ring variant): lane l of a warp step reads row wrow0 + c + l of the tile;
// c is a compile-time constant after unrolling, so it folds into the LDS immediate offset
__device__ __forceinline__ double qe_stage1_f64(const unsigned char *col, int wrow0, int c)
{
    return *(const double *)(col + (size_t)(wrow0 + (int)qe_lane()) * 8 + (size_t)c * 8);
}
__device__ __forceinline__ qe_i64 qe_stage1_i64(const unsigned char *col, int wrow0, int c)
{
    return *(const qe_i64 *)(col + (size_t)(wrow0 + (int)qe_lane()) * 8 + (size_t)c * 8);
}
__device__ __forceinline__ qe_i32 qe_stage1_i32(const unsigned char *col, int wrow0, int c)
{
    return *(const qe_i32 *)(col + (size_t)(wrow0 + (int)qe_lane()) * 4 + (size_t)c * 4);
}
__device__ __forceinline__ bool qe_stage1_bit(const unsigned char *bitmap, int wrow0, int c)
{
    return (*(const qe_u32 *)(bitmap + (size_t)(wrow0 >> 5) * 4 + (size_t)(c >> 5) * 4) >> qe_lane()) & 1u;
}

// the same through a 32-bit shared-window address (ring variant): the kernel keeps ONE opaque base in a
// register and every shared access is base + constant.  With C++ pointers to __shared__ objects nvcc
// re-derives each address inside the loop (MOV window base, VIADD offset, S2R SR_CgaCtaId, LEA): three S2R
// and ~30 address instructions per consumer warp and tile, measured as `wait` / short-scoreboard stalls.
__device__ __forceinline__ double qe_stage1_f64(qe_u32 col, int wrow0, int c)
{
    double v;
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(col + (qe_u32)(wrow0 + (int)qe_lane()) * 8u + (qe_u32)c * 8u));
    return v;
}
__device__ __forceinline__ qe_i64 qe_stage1_i64(qe_u32 col, int wrow0, int c)
{
    qe_i64 v;
    asm volatile("ld.shared.s64 %0, [%1];" : "=l"(v) : "r"(col + (qe_u32)(wrow0 + (int)qe_lane()) * 8u + (qe_u32)c * 8u));
    return v;
}
__device__ __forceinline__ qe_i32 qe_stage1_i32(qe_u32 col, int wrow0, int c)
{
    qe_i32 v;
    asm volatile("ld.shared.s32 %0, [%1];" : "=r"(v) : "r"(col + (qe_u32)(wrow0 + (int)qe_lane()) * 4u + (qe_u32)c * 4u));
    return v;
}
__device__ __forceinline__ bool qe_stage1_bit(qe_u32 bitmap, int wrow0, int c)
{
    qe_u32 w;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(w) : "r"(bitmap + (qe_u32)(wrow0 >> 5) * 4u + (qe_u32)(c >> 5) * 4u));
    return (w >> qe_lane()) & 1u;
}

// row r of a staged column, any r (the ring kernel's sparse path: one kept row per lane, qe_ring_scatter_list)
__device__ __forceinline__ double qe_stage_at_f64(qe_u32 col, int r)
{
    double v;
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(col + (qe_u32)r * 8u));
    return v;
}
__device__ __forceinline__ qe_i64 qe_stage_at_i64(qe_u32 col, int r)
{
    qe_i64 v;
    asm volatile("ld.shared.s64 %0, [%1];" : "=l"(v) : "r"(col + (qe_u32)r * 8u));
    return v;
}
__device__ __forceinline__ qe_i32 qe_stage_at_i32(qe_u32 col, int r)
{
    qe_i32 v;
    asm volatile("ld.shared.s32 %0, [%1];" : "=r"(v) : "r"(col + (qe_u32)r * 4u));
    return v;
}
__device__ __forceinline__ bool qe_stage_at_bit(qe_u32 bitmap, int r)
{
    qe_u32 w;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(w) : "r"(bitmap + ((qe_u32)r >> 5) * 4u));
    return (w >> ((qe_u32)r & 31u)) & 1u;
}

__device__ __forceinline__ qe_u32 qe_lds_u32(qe_u32 a)
{
    qe_u32 v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a) : "memory");
    return v;
}
// list entries of the ring kernel's sparse path: u8 when a consumer warp's slice has at most 256 rows, else u16
template <int W>
__device__ __forceinline__ qe_u32 qe_lds_entry(qe_u32 a)
{
    qe_u32 v;
    if (W == 1) asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(a) : "memory");
    else asm volatile("ld.shared.u16 %0, [%1];" : "=r"(v) : "r"(a) : "memory");
    return v;
}
template <int W>
__device__ __forceinline__ void qe_sts_entry_if(bool ok, qe_u32 a, qe_u32 v)
{
    if (W == 1) asm volatile("{ .reg .pred q; setp.ne.u32 q, %0, 0; @q st.shared.u8 [%1], %2; }" :: "r"((qe_u32)ok), "r"(a), "r"(v) : "memory");
    else asm volatile("{ .reg .pred q; setp.ne.u32 q, %0, 0; @q st.shared.u16 [%1], %2; }" :: "r"((qe_u32)ok), "r"(a), "r"(v) : "memory");
}
__device__ __forceinline__ qe_u64 qe_lds_u64(qe_u32 a)
{
    qe_u64 v;
    asm volatile("ld.shared.u64 %0, [%1];" : "=l"(v) : "r"(a) : "memory");
    return v;
}
__device__ __forceinline__ void qe_lds_u32x4(qe_u32 a, qe_u32 &x, qe_u32 &y, qe_u32 &z, qe_u32 &w)
{
    asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(x), "=r"(y), "=r"(z), "=r"(w) : "r"(a) : "memory");
}
__device__ __forceinline__ void qe_sts_u32(qe_u32 a, qe_u32 v)
{
    asm volatile("st.shared.u32 [%0], %1;" :: "r"(a), "r"(v) : "memory");
}
__device__ __forceinline__ void qe_sts_u64(qe_u32 a, qe_u64 v)
{
    asm volatile("st.shared.u64 [%0], %1;" :: "r"(a), "l"(v) : "memory");
}
__device__ __forceinline__ void qe_sts_u32x4(qe_u32 a, qe_u32 x, qe_u32 y, qe_u32 z, qe_u32 w)
{
    asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" :: "r"(a), "r"(x), "r"(y), "r"(z), "r"(w) : "memory");
}
__device__ __forceinline__ qe_u32 qe_opaque_u32(qe_u32 x)
{
    asm volatile("" : "+r"(x));
    return x;
}

// K1, staged variant.  Additional members of P:
//   static constexpr int NSTAGES, STAGE_BYTES (stride of one stage), STAGE_TX (bytes the copies of one stage move);
//   static void issue(const QeParams&, qe_i64 tile_row0, qe_u32 stage_smem, qe_u32 mbar);   one thread
//   static void load_stage(const unsigned char *stage, int r, Row (&rows)[VEC]);             r = row within the tile
template <class P>
__device__ __forceinline__ void qe_filter_staged_body(const QeParams &p)
{
    constexpr int STEPS = P::STEPS, VEC = P::VEC, NST = P::NSTAGES;
    constexpr int WARP_ROWS = STEPS * 32 * VEC;
    constexpr int TILE = QE_WARPS * WARP_ROWS;

    extern __shared__ __align__(128) unsigned char qe_dyn_smem[];
    __shared__ __align__(8) qe_u64 s_full[NST];
    __shared__ qe_u32 s_tile[NST];
    __shared__ qe_u32 s_warp_cnt[QE_WARPS];
    __shared__ qe_u64 s_lb_sum[QE_WARPS];
    __shared__ qe_u32 s_lb_found[QE_WARPS];

    const qe_u32 lane = qe_lane(), warp = qe_warp();

    // warp 1, lane 0 is the producer: it owns the tickets and issues every bulk copy
    const bool producer = threadIdx.x == 32;
    if (producer) {
#pragma unroll
        for (int st = 0; st < NST; st++) qe_mbar_init(qe_smem_addr(&s_full[st]), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
#pragma unroll
        for (int st = 0; st < NST; st++) {
            const qe_u32 t = atomicAdd(p.ticket, 1u);
            s_tile[st] = t;
            if (t < p.ntiles) {
                qe_mbar_expect_tx(qe_smem_addr(&s_full[st]), (qe_u32)P::STAGE_TX);
                P::issue(p, (qe_i64)t * TILE, qe_smem_addr(qe_dyn_smem + (size_t)st * P::STAGE_BYTES), qe_smem_addr(&s_full[st]));
            }
        }
    }
    __syncthreads();

    for (qe_u32 it = 0;; it++) {
        const int st = (int)(it % NST);
        const qe_u32 tile = s_tile[st];
        if (tile >= p.ntiles) break;                      // tickets only grow: every later stage is past the end too
        const qe_i64 warp_row0 = (qe_i64)tile * TILE + (qe_i64)warp * WARP_ROWS;
        const unsigned char *stage = qe_dyn_smem + (size_t)st * P::STAGE_BYTES;

        qe_mbar_wait(qe_smem_addr(&s_full[st]), (it / NST) & 1u);

        typename P::Out outs[STEPS][VEC];
        qe_u32 keep[STEPS];
        {
            typename P::Row rows[STEPS][VEC];
#pragma unroll
            for (int s = 0; s < STEPS; s++)
                P::load_stage(stage, (int)(warp * WARP_ROWS + s * 32 * VEC + lane * VEC), rows[s]);
            qe_rows_eval<P>(p, warp_row0, rows, outs, keep);
        }
        qe_u32 rank0[STEPS], step_off[STEPS];
        const qe_u32 wcnt = qe_warp_ranks<STEPS, VEC>(keep, rank0, step_off);
        if (lane == 0) s_warp_cnt[warp] = wcnt;
        __syncthreads();                                  // stage `st` and s_tile[st] have been consumed by everyone

        if (producer) {                                   // refill the stage for iteration it + NST
            const qe_u32 t = atomicAdd(p.ticket, 1u);
            s_tile[st] = t;
            if (t < p.ntiles) {
                qe_mbar_expect_tx(qe_smem_addr(&s_full[st]), (qe_u32)P::STAGE_TX);
                P::issue(p, (qe_i64)t * TILE, qe_smem_addr(stage), qe_smem_addr(&s_full[st]));
            }
        }
        qe_u32 woff = 0, tile_total = 0;
#pragma unroll
        for (int w = 0; w < QE_WARPS; w++) {
            const qe_u32 c = s_warp_cnt[w];
            woff += (qe_u32)w < warp ? c : 0u;
            tile_total += c;
        }
        const qe_u64 excl = (qe_out_base_full(p) & p.out_base_mask) + qe_block_lookback(p.tile_state, tile, tile_total, s_lb_sum, s_lb_found);
        if (threadIdx.x == 0 && tile == p.ntiles - 1) *p.out_count = (qe_out_base_full(p) & ~p.out_base_mask) + excl + tile_total;
        qe_warp_scatter<P>(p, warp_row0, excl + woff, outs, keep, rank0, step_off);
    }
}

// ---- K1, ring variant (the default) ------------------------------------------------
//
// One persistent CTA per SM, warp-specialised around a ring of NSTAGES shared-memory
// stages (a stage = one tile of every referenced column, 16-48 KB):
//
//   warp 0      producer   claims tiles through the ticket counter and issues one
//                          cp.async.bulk per column run into the next free stage
//   warp 1      publisher  when a tile has been evaluated: sums the per-warp kept-row
//                          counts and publishes the tile's descriptor (count + 1)
//   warps 2..   look-back  NLB warps: exclusive prefix of each of this CTA's tiles (round robin)
//   then        consumers  phase A (evaluate the filter on a tile from shared memory,
//                          ballot -> keep-mask words + per-warp counts) and, DEFER
//                          tiles later, phase C (ordered scatter of the kept rows:
//                          selection vector + compacted outputs)
//
// Why this shape (measured on B200): stable compaction cannot write
// tile t before every earlier tile has been counted, and on a 148-SM part a
// descriptor needs an L2 round trip (~500-900 cycles under streaming load) to travel
// between SMs while a 32 KB tile streams in ~1 500 cycles.  A kernel that holds its
// tile in registers while it waits stalls the loads; here the tile waits in shared
// memory, the bulk copies of the next tiles keep running (bytes in flight are bounded
// by the 227 KB of shared memory, not by registers), and the consumers always have a
// phase A to run while a prefix is outstanding.
//
// Look-back over a persistent grid: tickets are handed out in increasing order, so
// between two consecutive tiles t' < t of one CTA lie only tiles of other CTAs, and
//     excl(t) = incl(t') + sum of the counts of tiles t'+1 .. t-1.
// The look-back warp therefore never hunts for a predecessor with a known prefix: it
// reads that gap (on average gridDim.x - 1 descriptors, all in flight at once) and
// adds it to its own running total.  Every predecessor was claimed earlier by a
// resident CTA, so the spin cannot deadlock.
//
// All hand-offs are mbarriers (arrive = release, try_wait = acquire at CTA scope):
//   full[st]   producer arrive + bulk-copy bytes      -> consumers
//   claim[st]  producer (tile id known)               -> publisher, look-back
//   eval[st]   NCW consumer warps (phase A done)      -> publisher, look-back, phase C
//   pref[st]   look-back (s_excl[st] written)         -> phase C
//   empty[st]  NCW consumers + publisher + look-back  -> producer
//
// Additional members of P:
//   static constexpr int TILE_ROWS, NCW, NLB, NSTAGES, DEFER, LOADS, STAGE_BYTES, STAGE_TX;
//   static void issue(const QeParams&, qe_i64 tile_row0, qe_u32 stage_smem, qe_u32 mbar, qe_u64 l2_policy);   one thread
//   static void load_row(const unsigned char *stage, int wrow0, int c, Row&);   row wrow0 + c + lane of the tile (wrow0, c: multiples of 32)
//   struct Ptrs; static Ptrs out_ptrs(const QeParams&, qe_u64 base);          value output arrays, advanced to position base
//   static void emit_ptr(const Ptrs&, qe_u32 idx, const Out&, bool ok);       value outputs at position base + idx when ok

__device__ __forceinline__ void qe_mbar_arrive(qe_u32 mbar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(mbar) : "memory");
}

__device__ __forceinline__ qe_u32 qe_ld_relaxed_u32(const qe_u32 *p)
{
    qe_u32 v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

__device__ __forceinline__ void qe_st_relaxed_u32(qe_u32 *p, qe_u32 v)
{
    asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}

// Predicated stores.  Written as inline PTX so that phase C stays branch-free: left to
// itself the compiler wraps every warp step in a divergent region (BSSY / BSYNC) and
// re-derives the addresses inside it.
__device__ __forceinline__ void qe_stg_if(bool ok, double *a, double v)
{
    asm volatile("{ .reg .pred q; setp.ne.u32 q, %0, 0; @q st.global.f64 [%1], %2; }" :: "r"((qe_u32)ok), "l"(a), "d"(v) : "memory");
}
__device__ __forceinline__ void qe_stg_if(bool ok, qe_i64 *a, qe_i64 v)
{
    asm volatile("{ .reg .pred q; setp.ne.u32 q, %0, 0; @q st.global.s64 [%1], %2; }" :: "r"((qe_u32)ok), "l"(a), "l"(v) : "memory");
}
__device__ __forceinline__ void qe_stg_if(bool ok, qe_i32 *a, qe_i32 v)
{
    asm volatile("{ .reg .pred q; setp.ne.u32 q, %0, 0; @q st.global.s32 [%1], %2; }" :: "r"((qe_u32)ok), "l"(a), "r"(v) : "memory");
}
__device__ __forceinline__ void qe_stg_if(bool ok, qe_u32 *a, qe_u32 v)
{
    asm volatile("{ .reg .pred q; setp.ne.u32 q, %0, 0; @q st.global.u32 [%1], %2; }" :: "r"((qe_u32)ok), "l"(a), "r"(v) : "memory");
}
// keeps a once-per-tile pointer in its register pair (the optimiser otherwise folds the 64-bit
// base back into every address computation: 6 instructions per store instead of 1)
template <class T>
__device__ __forceinline__ T *qe_opaque_ptr(T *q)
{
    asm volatile("" : "+l"(q));
    return q;
}

// Phase C of the ring kernel for one warp: kept rows of its slice, in row order, to
// consecutive output positions starting at wbase.  One store instruction covers one
// warp step = one contiguous run of the outputs.  Two code paths, chosen per warp and
// tile by the number of kept rows: a fully unrolled branch-free one (predicated
// stores), and for sparse tiles a loop over the steps that kept anything.
// Compacted bit columns of one consumer warp's slice of a tile: the kept rows' bits are appended step by step to a
// 64-bit window per bit column (uniform over the warp; one REDUX.OR per step and column), and every time 32 bits are
// complete the low word is OR-ed into the pre-zeroed bitmap by lane 0.  (Round 1 shifted and OR-ed two words per step and
// column, qe_store_bits_compact: ~20 instructions and two atomics, against ~8 and one atomic per 32 kept rows.)
template <int NB>
struct QeBitRun {
    qe_u64 acc[NB > 0 ? NB : 1];
    qe_u64 word;              // index of the bitmap word the low half of acc belongs to
    qe_u32 fill;              // bits of acc in use
};

template <class P>
__device__ __forceinline__ void qe_bit_run_flush(const QeParams &p, QeBitRun<P::NBITCOLS> &br)
{
    if (qe_lane() == 0) {
#pragma unroll
        for (int b = 0; b < P::NBITCOLS; b++) {
            qe_u32 *bm = P::bitcol(p, b);
            const qe_u32 lo = (qe_u32)br.acc[b];
            if (bm && lo) atomicOr(bm + br.word, lo);
        }
    }
}

template <class P, bool CHECK_CAP, bool HAS_SEL>
__device__ __forceinline__ void qe_ring_scatter_step(const QeParams &p, const typename P::Ptrs &ptrs, qe_u32 *selp,
                                                     qe_u32 stage, int wrow0, int c, qe_u32 row_id, qe_u64 wbase,
                                                     qe_u32 cap_left, qe_u32 m, qe_u32 &off, QeBitRun<P::NBITCOLS> &br)
{
    const qe_u32 lane = qe_lane();
    const bool kept = (m >> lane) & 1u;
    const qe_u32 idx = off + __popc(m & qe_lanemask_lt());
    typename P::Row row;
    typename P::Out o;
    P::load_row(stage, wrow0, c, row);
    (void)P::eval(p, row, o);
    const bool ok = kept && (!CHECK_CAP || idx < cap_left);
    if (HAS_SEL) qe_stg_if(ok, selp + idx, row_id);
    P::emit_ptr(ptrs, idx, o, ok);
    if (P::NBITCOLS > 0) {
        (void)wbase;
        const qe_u32 mine = 1u << (idx - off);                    // this row's place among the step's kept rows
#pragma unroll
        for (int b = 0; b < P::NBITCOLS; b++) {
            const qe_u32 w = __reduce_or_sync(QE_FULL_MASK, (ok && P::bit(o, b)) ? mine : 0u);   // rows past the capacity leave no bit
            br.acc[b] |= (qe_u64)w << br.fill;
        }
        br.fill += (qe_u32)__popc(m);
        if (br.fill >= 32u) {                                     // uniform
            qe_bit_run_flush<P>(p, br);
#pragma unroll
            for (int b = 0; b < P::NBITCOLS; b++) br.acc[b] >>= 32;
            br.fill -= 32u;
            br.word += 1ull;
        }
    }
    off += __popc(m);
}

template <class P, int STEPS, bool CHECK_CAP, bool HAS_SEL>
__device__ __forceinline__ void qe_ring_scatter(const QeParams &p, qe_u32 stage, int wrow0, qe_u32 first_row,
                                                qe_u64 wbase, qe_u32 masks, qe_u32 cnt)
{
    // stage, masks: shared-window addresses of the tile's stage and of this warp's STEPS keep-mask words
    const qe_u32 lane = qe_lane();
    const typename P::Ptrs ptrs = P::out_ptrs(p, wbase);
    qe_u32 *const selp = HAS_SEL ? qe_opaque_ptr(p.sel + wbase) : (qe_u32 *)0;
    qe_u32 cap_left = 0xffffffffu;                            // outputs of this warp that still fit (CHECK_CAP only)
    if (CHECK_CAP) {
        const qe_u64 cap = (qe_u64)p.out_capacity;
        cap_left = wbase >= cap ? 0u : (cap - wbase > 0xffffffffull ? 0xffffffffu : (qe_u32)(cap - wbase));
    }
    qe_u32 off = 0;                                           // kept rows of this warp in earlier steps
    QeBitRun<P::NBITCOLS> br;
#pragma unroll
    for (int b = 0; b < (P::NBITCOLS > 0 ? P::NBITCOLS : 1); b++) br.acc[b] = 0ull;
    br.word = wbase >> 5;
    br.fill = (qe_u32)wbase & 31u;
    // Which path is decided by the number of warp steps that kept anything, not by the row count: the loop
    // costs ~27 instructions per non-empty step, the unrolled path ~15 per step.  (Deciding by rows kept,
    // dense from 25 % up, left 5-20 % selectivity in the loop with every step non-empty: 2.2-2.55 ms per
    // 10^9 rows where 1 % took 1.45 and 30 % 1.88.)
    const qe_u32 mine = lane < (qe_u32)STEPS ? qe_lds_u32(masks + lane * 4u) : 0u;
    qe_u32 act = __ballot_sync(QE_FULL_MASK, mine != 0u);
    (void)cnt;
    if ((qe_u32)__popc(act) * 2u > (qe_u32)STEPS) {
        qe_u32 mk[STEPS];
        if (STEPS % 4 == 0) {
#pragma unroll
            for (int s = 0; s < STEPS; s += 4) qe_lds_u32x4(masks + (qe_u32)s * 4u, mk[s], mk[s + 1], mk[s + 2], mk[s + 3]);
        } else {
#pragma unroll
            for (int s = 0; s < STEPS; s++) mk[s] = qe_lds_u32(masks + (qe_u32)s * 4u);
        }
#pragma unroll
        for (int s = 0; s < STEPS; s++)
            qe_ring_scatter_step<P, CHECK_CAP, HAS_SEL>(p, ptrs, selp, stage, wrow0, s * 32, first_row + (qe_u32)(s * 32), wbase,
                                                        cap_left, mk[s], off, br);
    } else {
        // Sparse tiles: one non-empty step per iteration.  (Two steps per iteration -- both rows loaded and evaluated
        // before either is stored, to break the serial shuffle -> load -> evaluate -> rank -> store chain -- measured
        // SLOWER at 1 % selectivity, 1.36 against 1.32 ms per 10^9 rows, and equal at 5 %: most warps meet a single
        // non-empty step per tile and pay for the empty second one; profiles/r02_sweep_01_sparse_path.jsonl.)
        while (act) {
            const int s = __ffs((int)act) - 1;
            act &= act - 1u;
            const qe_u32 m = __shfl_sync(QE_FULL_MASK, mine, s);
            qe_ring_scatter_step<P, CHECK_CAP, HAS_SEL>(p, ptrs, selp, stage, wrow0, s * 32, first_row + (qe_u32)(s * 32), wbase,
                                                        cap_left, m, off, br);
        }
    }
    if (P::NBITCOLS > 0 && br.fill != 0u) qe_bit_run_flush<P>(p, br);     // the slice's last, partial word
}

// Phase C for a warp whose slice kept at most P::LIST_CAP rows of the tile: phase A left their row numbers (within the
// warp's slice, ascending) as a packed list in shared memory, so lane l stores entry l -- one pass of coalesced stores
// whatever steps the rows came from, where the mask paths above walk the steps (~27 instructions per non-empty step).
// Only for pipelines without compacted bit columns, and only when every output of the tile fits the capacity.
template <class P, bool HAS_SEL>
__device__ __forceinline__ void qe_ring_scatter_list(const QeParams &p, qe_u32 stage, int wrow0, qe_u32 first_row_of_slice,
                                                     qe_u64 wbase, qe_u32 list, qe_u32 cnt)
{
    constexpr int LW = P::TILE_ROWS / P::NCW <= 256 ? 1 : 2;  // bytes per list entry
    const qe_u32 lane = qe_lane();
    const typename P::Ptrs ptrs = P::out_ptrs(p, wbase);
    qe_u32 *const selp = HAS_SEL ? qe_opaque_ptr(p.sel + wbase) : (qe_u32 *)0;
#pragma unroll
    for (qe_u32 base = 0; base < (qe_u32)P::LIST_CAP; base += 32u) {
        if (base >= cnt) break;
        const qe_u32 i = base + lane;
        const bool ok = i < cnt;
        const qe_u32 r = qe_lds_entry<LW>(list + (qe_u32)LW * (ok ? i : 0u));
        typename P::Row row;
        typename P::Out o;
        P::load_at(stage, wrow0 + (int)r, row);
        (void)P::eval(p, row, o);
        if (HAS_SEL) qe_stg_if(ok, selp + i, first_row_of_slice + r);
        P::emit_ptr(ptrs, i, o, ok);
    }
}

// bulk copy with an L2 evict-first policy: scanned columns are read exactly once, so they should not
// push the (re-read) tile descriptors or the partially written output lines out of L2
__device__ __forceinline__ void qe_bulk_g2s_stream(qe_u32 dst_smem, const void *src, qe_u32 bytes, qe_u32 mbar, qe_u64 policy)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
                 :: "r"(dst_smem), "l"(src), "r"(bytes), "r"(mbar), "l"(policy) : "memory");
}

__device__ __forceinline__ qe_u64 qe_policy_evict_first()
{
    qe_u64 pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}

__device__ __forceinline__ void qe_prefetch_l2_keep(const void *p)
{
    asm volatile("prefetch.global.L2::evict_last [%0];" :: "l"(p));
}

// Bits [lo, hi) of a pre-zeroed bitmap to one, by one warp: whole words with plain stores (no other run reaches into
// them), the two end words with an OR (neighbouring runs -- other tiles, other CTAs -- meet there).
__device__ __forceinline__ void qe_set_bit_run(qe_u32 *bitmap, qe_u64 lo, qe_u64 hi)
{
    if (hi <= lo) return;
    const qe_u64 w0 = lo >> 5, w1 = (hi - 1ull) >> 5;
    for (qe_u64 w = w0 + qe_lane(); w <= w1; w += 32ull) {
        qe_u32 m = 0xffffffffu;
        if (w == w0) m &= 0xffffffffu << ((qe_u32)lo & 31u);
        if (w == w1) m &= 0xffffffffu >> (31u - ((qe_u32)(hi - 1ull) & 31u));
        if (w == w0 || w == w1) atomicOr(bitmap + w, m);
        else bitmap[w] = m;
    }
}

#define QE_RING_LB_PER_LANE 6        // descriptors per lane per look-back round trip (192 per warp)

// phase timing of the ring kernel (QE_DEBUG_FLAGS & 16; tools/phase_timing.py): cycles per CTA in p.debug[cta][16]
//  0 producer: wait empty   1 producer: wait load throttle   2 producer: tiles
//  3 look-back warp 0: wait eval   4 first descriptor read   5 re-poll cycles   6 re-polls   7 wait previous prefix   8 gap length
//  9 consumer warp 0: wait full   10 phase A   11 wait prefix   12 phase C   14 look-back warp 0: first reads that came too early
#define QE_RT_BEGIN() qe_u64 rt_t0_ = (QE_DEBUG_FLAGS & 16) ? clock64() : 0
#define QE_RT_ADD(i) do { if (QE_DEBUG_FLAGS & 16) { const qe_u64 n_ = clock64(); rt_acc[i] += n_ - rt_t0_; rt_t0_ = n_; } } while (0)
#define QE_RT_RESET() do { if (QE_DEBUG_FLAGS & 16) rt_t0_ = clock64(); } while (0)
#define QE_RT_FLUSH(lo, hi) do { if ((QE_DEBUG_FLAGS & 16) && lane == 0 && blockIdx.x < 2048) { \
        for (int i_ = (lo); i_ <= (hi); i_++) p.debug[blockIdx.x * 16 + i_] = rt_acc[i_]; } } while (0)

template <class P>
__device__ __forceinline__ void qe_filter_ring_body(const QeParams &p)
{
    constexpr int NST = P::NSTAGES, NCW = P::NCW, TILE = P::TILE_ROWS, DEFER = P::DEFER, NLB = P::NLB;
    constexpr int WROWS = TILE / NCW, STEPS = WROWS / 32;
    static_assert(WROWS % 32 == 0 && NCW <= 32 && DEFER >= 1 && DEFER < NST && NLB >= 1 && NLB <= NST, "ring shape");

    // control block, addressed as (one opaque 32-bit shared-window base) + constant: see qe_stage1_*(qe_u32, ...)
    constexpr qe_u32 O_FULL = 0, O_CLAIM = 8 * NST, O_EVAL = 16 * NST, O_PREF = 24 * NST, O_EMPTY = 32 * NST;   // mbarriers [NST]
    constexpr qe_u32 O_INCL = 40 * NST;                       // u64 [NST]: kept rows up to and including the tile
    constexpr qe_u32 O_WBASE = 48 * NST;                      // u64 [NST][NCW]: output position of each consumer warp's first kept row
    constexpr qe_u32 O_TILE = O_WBASE + 8 * NST * NCW;        // u32 [NST] each: tile, lo, row0, fits
    constexpr qe_u32 O_LO = O_TILE + 4 * NST, O_ROW0 = O_LO + 4 * NST, O_FITS = O_ROW0 + 4 * NST;
    constexpr qe_u32 O_CNT = (O_FITS + 4 * NST + 15u) & ~15u; // u32 [NST][NCW]
    constexpr qe_u32 O_MASK = (O_CNT + 4 * NST * NCW + 15u) & ~15u;   // u32 [NST][TILE / 32]
    // sparse tiles (qe_ring_scatter_list, sweepers): u8 / u16 [NST][NCW][LIST_CAP] row numbers within the warp's slice
    constexpr bool LIST = P::LIST_CAP > 0 && (P::NBITCOLS == 0 || P::NSW > 0);   // compacted bit columns: through the sweepers only
    constexpr qe_u32 LIST_CAP = LIST ? (qe_u32)P::LIST_CAP : 0u;
    constexpr int NSW = LIST ? P::NSW : 0;                  // sweeper warps (0: a consumer warp stores its own list in phase C)
    constexpr qe_u32 O_LIST = (O_MASK + 4 * NST * (TILE / 32) + 15u) & ~15u;
    constexpr int LW = WROWS <= 256 ? 1 : 2;                  // bytes per list entry
    constexpr qe_u32 CTL_BYTES = O_LIST + (qe_u32)LW * NST * NCW * LIST_CAP;

    extern __shared__ __align__(128) unsigned char qe_dyn_smem[];
    __shared__ __align__(16) unsigned char s_ctl[CTL_BYTES];

    const qe_u32 lane = qe_lane(), warp = qe_warp();
    qe_u32 *const desc = (qe_u32 *)p.tile_state;             // one 32-bit descriptor per tile: 0 = not counted yet, else count + 1
    const qe_u32 cb = qe_opaque_u32(qe_smem_addr(s_ctl));
    const qe_u32 dyn = qe_opaque_u32(qe_smem_addr(qe_dyn_smem));

    if (threadIdx.x == 0) {
#pragma unroll
        for (int st = 0; st < NST; st++) {
            qe_mbar_init(cb + O_FULL + 8u * st, 1);
            qe_mbar_init(cb + O_CLAIM + 8u * st, 1);
            qe_mbar_init(cb + O_EVAL + 8u * st, NCW);
            qe_mbar_init(cb + O_PREF + 8u * st, 1);
            qe_mbar_init(cb + O_EMPTY + 8u * st, NCW + 2 + NSW);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    // every role walks the stages round robin: st = it % NST and par = (it / NST) & 1 are carried, not divided out
#define QE_RING_NEXT(st_, par_, by_) do { st_ += (by_); if (st_ >= (qe_u32)NST) { st_ -= (qe_u32)NST; par_ ^= 1u; } } while (0)

    if (warp == 0) {
        // ------------------------------------------------------------ producer
        if (lane == 0) {
            const qe_u64 policy = qe_policy_evict_first();
            qe_u32 t = atomicAdd(p.ticket, 1u);
            qe_u32 lo = 0, ended = 0;
            qe_u64 rt_acc[16] = {};
            QE_RT_BEGIN();
            for (qe_u32 it = 0;; it++) {
                const qe_u32 st = it % NST;
                QE_RT_RESET();
                if (it >= NST) qe_mbar_wait(cb + O_EMPTY + 8u * st, ((it / NST) - 1u) & 1u);
                QE_RT_ADD(0);
                // at most LOADS tiles in flight: more only queue up in the memory system, and every
                // descriptor read of the look-back then waits behind them
                if (P::LOADS > 0 && it >= (qe_u32)P::LOADS && ended == 0u) {
                    const qe_u32 jt = it - (qe_u32)P::LOADS;
                    qe_mbar_wait(cb + O_FULL + 8u * (jt % NST), (jt / NST) & 1u);
                }
                QE_RT_ADD(1);
                rt_acc[2] += 1;
                qe_sts_u32(cb + O_TILE + 4u * st, t);
                qe_sts_u32(cb + O_LO + 4u * st, lo);           // tiles [lo, t) belong to other CTAs
                qe_sts_u32(cb + O_ROW0 + 4u * st, p.sel_base + t * (qe_u32)TILE);   // selection-vector value of the tile's first row
                lo = t + 1u;
                qe_mbar_arrive(cb + O_CLAIM + 8u * st);
                if (t >= p.ntiles) {
                    // end marker: wakes the consumers (first one only), carries no bytes; every look-back
                    // warp must meet one on its own round-robin slot, so NLB of them are written
                    if (ended == 0u) qe_mbar_arrive(cb + O_FULL + 8u * st);
                    if (++ended == (qe_u32)NLB) { QE_RT_FLUSH(0, 2); break; }
                    continue;
                }
                // the tile's descriptor line should be in L2 when it is published and read, NST tiles from now
                if (!(QE_DEBUG_FLAGS & 32)) qe_prefetch_l2_keep(desc + t);
                const qe_u32 full = cb + O_FULL + 8u * st;
                qe_mbar_expect_tx(full, (qe_u32)P::STAGE_TX);
                P::issue(p, (qe_i64)t * TILE, dyn + st * (qe_u32)P::STAGE_BYTES, full, policy);
                t = atomicAdd(p.ticket, 1u);                  // next ticket: its round trip overlaps the wait for a free stage
            }
        }
    } else if (warp == 1) {
        // ----------------------------------------------------------- publisher
        qe_u32 st = 0, par = 0;
        for (;;) {
            qe_mbar_wait(cb + O_CLAIM + 8u * st, par);
            const qe_u32 t = qe_lds_u32(cb + O_TILE + 4u * st);
            if (t >= p.ntiles) break;
            qe_mbar_wait(cb + O_EVAL + 8u * st, par);
            qe_u32 c = lane < (qe_u32)NCW ? qe_lds_u32(cb + O_CNT + 4u * ((qe_u32)NCW * st + lane)) : 0u;
            c = __reduce_add_sync(QE_FULL_MASK, c);
            if (lane == 0) {
                qe_st_relaxed_u32(desc + t, c + 1u);
                qe_mbar_arrive(cb + O_EMPTY + 8u * st);
            }
            QE_RING_NEXT(st, par, 1u);
        }
    } else if (warp < 2u + (qe_u32)NLB) {
        // ----------------------------------------------------------- look-back
        // NLB warps take this CTA's tiles round robin.  A prefix costs one L2 round trip after the
        // tile's count is visible grid-wide (~1.4 us under streaming load, measured), a 16-32 KB tile
        // streams in 0.4-0.7 us, so several must be in flight; the gap sums are independent, only the
        // final add chains through shared memory (the inclusive count of the previous tile).
        qe_u64 rt_acc[16] = {};
        QE_RT_BEGIN();
        const qe_u64 base_full = qe_out_base_full(p);          // append mode: rows of earlier batches
        const qe_u64 base0 = base_full & p.out_base_mask;
        for (qe_u32 it = warp - 2u;; it += (qe_u32)NLB) {
            const qe_u32 st = it % NST;
            const qe_u32 par = (it / NST) & 1u;
            qe_mbar_wait(cb + O_CLAIM + 8u * st, par);
            const qe_u32 t = qe_lds_u32(cb + O_TILE + 4u * st);
            if (t >= p.ntiles) { if (warp == 2) { QE_RT_FLUSH(3, 8); QE_RT_FLUSH(14, 14); } break; }
            const qe_u32 lo = qe_lds_u32(cb + O_LO + 4u * st);
            QE_RT_RESET();
            // Wait for this CTA's own count first: the grid moves in step, so by then the tiles of the
            // other CTAs in the gap are counted (or about to be) and the descriptors are read once
            // instead of being polled for a whole tile time (148 warps polling five cache lines slowed
            // the very stores they were waiting for: measured 5-12 us per prefix).
            qe_mbar_wait(cb + O_EVAL + 8u * st, par);
            QE_RT_ADD(3);
            rt_acc[8] += t - lo;
            const qe_u32 mine = lane < (qe_u32)NCW ? qe_lds_u32(cb + O_CNT + 4u * ((qe_u32)NCW * st + lane)) : 0u;
            const qe_u32 c = __reduce_add_sync(QE_FULL_MASK, mine);
            // the gap [lo, t): tiles of other CTAs claimed between this CTA's previous tile and this one
            qe_u32 gap = 0;
            for (qe_u32 base = lo; base < t; base += 32u * QE_RING_LB_PER_LANE) {
                qe_u32 d[QE_RING_LB_PER_LANE];
#pragma unroll
                for (int k = 0; k < QE_RING_LB_PER_LANE; k++) {   // every read in flight before the first is used
                    const qe_u32 idx = base + (qe_u32)k * 32u + lane;
                    d[k] = idx < t ? qe_ld_relaxed_u32(desc + idx) : 1u;
                }
                if (QE_DEBUG_FLAGS & 16) {                        // time of the first read = one L2 round trip
                    bool inval = false;
#pragma unroll
                    for (int k = 0; k < QE_RING_LB_PER_LANE; k++) inval |= d[k] == 0u;
                    if (__any_sync(QE_FULL_MASK, inval)) rt_acc[14] += 1;      // first read came too early
                    QE_RT_ADD(4);
                }
                // Re-read what was not published yet: every missing descriptor again in one go, the warp
                // stays converged.  No __nanosleep here: on B200 the shortest sleep measured ~1 us (ncu:
                // two sleeps per tile = the whole 2.3 us tile time), a re-read is its own ~0.5 us back-off.
                for (;;) {
                    bool pending = false;
#pragma unroll
                    for (int k = 0; k < QE_RING_LB_PER_LANE; k++) pending |= d[k] == 0u;
                    if (!__any_sync(QE_FULL_MASK, pending)) break;
                    if (QE_DEBUG_FLAGS & 16) rt_acc[6] += 1;
#pragma unroll
                    for (int k = 0; k < QE_RING_LB_PER_LANE; k++) {
                        const qe_u32 idx = base + (qe_u32)k * 32u + lane;
                        if (d[k] == 0u) d[k] = qe_ld_relaxed_u32(desc + idx);
                    }
                }
#pragma unroll
                for (int k = 0; k < QE_RING_LB_PER_LANE; k++) gap += d[k] - 1u;
                QE_RT_ADD(5);
            }
            gap = __reduce_add_sync(QE_FULL_MASK, gap);
            qe_u64 before = base0;                            // kept rows up to and including this CTA's previous tile
            if (it > 0) {
                const qe_u32 jt = it - 1u;
                qe_mbar_wait(cb + O_PREF + 8u * (jt % NST), (jt / NST) & 1u);
                before = qe_lds_u64(cb + O_INCL + 8u * (jt % NST));
            }
            QE_RT_ADD(7);
            const qe_u64 excl = before + gap;
            // exclusive scan of the per-warp counts: every consumer warp gets its own base
            qe_u32 wofs = mine;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const qe_u32 v = __shfl_up_sync(QE_FULL_MASK, wofs, o);
                if (lane >= (qe_u32)o) wofs += v;
            }
            if (lane < (qe_u32)NCW) qe_sts_u64(cb + O_WBASE + 8u * ((qe_u32)NCW * st + lane), excl + (wofs - mine));
            __syncwarp();
            if (lane == 0) {
                qe_sts_u32(cb + O_FITS + 4u * st, excl + c <= (qe_u64)p.out_capacity ? 1u : 0u);
                qe_sts_u64(cb + O_INCL + 8u * st, excl + c);
                qe_mbar_arrive(cb + O_PREF + 8u * st);
                qe_mbar_arrive(cb + O_EMPTY + 8u * st);
                if (t == p.ntiles - 1u) *p.out_count = (base_full & ~p.out_base_mask) + excl + c;   // stays the running total
            }
            if (P::NIMPLIED > 0 && c != 0u) {
                // validity bitmaps the filter has decided (P::implied_col): this tile's rows are ones [excl, excl + c), as
                // far as they fit.  After the prefix is out, so nobody waits for it.
                const qe_u64 cap = (qe_u64)p.out_capacity;
                const qe_u64 lo = excl < cap ? excl : cap, hi = excl + c < cap ? excl + c : cap;
#pragma unroll
                for (int i = 0; i < P::NIMPLIED; i++) {
                    qe_u32 *bm = P::implied_col(p, i);
                    if (bm) qe_set_bit_run(bm, lo, hi);
                }
            }
        }
    } else if (NSW > 0 && warp < 2u + (qe_u32)NLB + (qe_u32)NSW) {
        // ----------------------------------------------------------- sweepers
        // Store the listed rows of every consumer warp that kept at most LIST_CAP rows of a tile, as soon as the tile's
        // prefix is known.  Measured on B200 (profiles/r02_sweep_02_sweepers.jsonl): a consumer warp that never runs
        // phase C and never waits for a prefix streams a 1 %-selective filter at the pure-read rate (1.19 ms per 10^9
        // rows, 1.22 with the lists built), against 1.31 ms with phase C in the consumers.  Thread q of the 32 * NSW
        // sweeper threads takes entries j0, j0 + JS, ... of consumer warp q % NCW.
        constexpr qe_u32 JS = (32u * (qe_u32)NSW) / (qe_u32)NCW;
        const qe_u32 q = (warp - (2u + (qe_u32)NLB)) * 32u + lane;
        const qe_u32 l = q % (qe_u32)NCW, j0 = q / (qe_u32)NCW;
        qe_u32 st = 0, par = 0;
        for (;;) {
            qe_mbar_wait(cb + O_CLAIM + 8u * st, par);
            const qe_u32 t = qe_lds_u32(cb + O_TILE + 4u * st);
            if (t >= p.ntiles) break;
            qe_mbar_wait(cb + O_EVAL + 8u * st, par);                          // counts and lists of every consumer warp
            const qe_u32 c = qe_lds_u32(cb + O_CNT + 4u * ((qe_u32)NCW * st + l));
            const bool mine = j0 < JS && j0 < c && c <= LIST_CAP;
            if (QE_DEBUG_FLAGS & 64) { __syncwarp(); if (lane == 0) qe_mbar_arrive(cb + O_EMPTY + 8u * st); }   // timing only: the stage is released before the prefix is known (the rows are then read from a stage that may be refilled)
            qe_mbar_wait(cb + O_PREF + 8u * st, par);                          // bases
            if (mine) {
                const qe_u64 wbase = qe_lds_u64(cb + O_WBASE + 8u * ((qe_u32)NCW * st + l));
                qe_u32 room = 0xffffffffu;                                     // outputs of this consumer warp that still fit
                if (qe_lds_u32(cb + O_FITS + 4u * st) == 0u) {
                    const qe_u64 cap = (qe_u64)p.out_capacity;
                    room = wbase >= cap ? 0u : (cap - wbase > 0xffffffffull ? 0xffffffffu : (qe_u32)(cap - wbase));
                }
                const qe_u32 stage = dyn + st * (qe_u32)P::STAGE_BYTES;
                const qe_u32 la = cb + O_LIST + (qe_u32)LW * LIST_CAP * ((qe_u32)NCW * st + l);
                const int slice0 = (int)l * WROWS;
                const qe_u32 first_row = qe_lds_u32(cb + O_ROW0 + 4u * st) + (qe_u32)slice0;
                const typename P::Ptrs ptrs = P::out_ptrs(p, wbase);
                qe_u32 *const selp = p.sel ? p.sel + wbase : (qe_u32 *)0;
                const qe_u32 end = c < room ? c : room;
                // two entries per round: both list reads, then both rows, then the stores -- a round is a chain of
                // shared-memory and store latencies, and a thread has up to LIST_CAP / JS entries of a dense-ish tile
                for (qe_u32 j = j0; j < end; j += 2u * JS) {
                    const qe_u32 j1 = j + JS;
                    const bool two = j1 < end;
                    const qe_u32 r0 = qe_lds_entry<LW>(la + (qe_u32)LW * j);
                    const qe_u32 r1 = qe_lds_entry<LW>(la + (qe_u32)LW * (two ? j1 : j));
                    typename P::Row row0, row1;
                    typename P::Out o0, o1;
                    P::load_at(stage, slice0 + (int)r0, row0);
                    P::load_at(stage, slice0 + (int)r1, row1);
                    (void)P::eval(p, row0, o0);
                    (void)P::eval(p, row1, o1);
                    if (selp) {
                        selp[j] = first_row + r0;
                        qe_stg_if(two, selp + j1, first_row + r1);
                    }
                    P::emit_ptr(ptrs, j, o0, true);
                    P::emit_ptr(ptrs, j1, o1, two);
                    if (P::NBITCOLS > 0) {
                        // compacted bit columns (validity, BOOLEAN values): one OR per set bit into the pre-zeroed bitmap --
                        // a listed tile keeps a handful of rows per consumer warp, and runs of different warps, tiles and
                        // CTAs meet inside a word anyway (qe_store_bits_compact)
                        const qe_u64 g0 = wbase + j, g1 = wbase + j1;
#pragma unroll
                        for (int b = 0; b < P::NBITCOLS; b++) {
                            qe_u32 *bm = P::bitcol(p, b);
                            if (bm) {
                                if (P::bit(o0, b)) atomicOr(bm + (g0 >> 5), 1u << ((qe_u32)g0 & 31u));
                                if (two && P::bit(o1, b)) atomicOr(bm + (g1 >> 5), 1u << ((qe_u32)g1 & 31u));
                            }
                        }
                    }
                }
            }
            __syncwarp();
            if (lane == 0 && !(QE_DEBUG_FLAGS & 64)) qe_mbar_arrive(cb + O_EMPTY + 8u * st);
            QE_RING_NEXT(st, par, 1u);
        }
    } else {
        // ----------------------------------------------------------- consumers
        const qe_u32 cw = warp - (2u + (qe_u32)NLB + (qe_u32)NSW);
        const int wrow0 = (int)cw * WROWS;                    // this warp's slice of every tile
        int n_valid = 0x7fffffff;                             // iterations that carry a tile (known once the end marker is seen)
        qe_u32 st_a = 0, par_a = 0;                           // stage / parity of phase A's tile (iteration it)
        qe_u32 st_c = 0, par_c = 0;                           // ... of phase C's tile (iteration it - DEFER)
        qe_u64 rt_acc[16] = {};
        QE_RT_BEGIN();
        for (int it = 0;; it++) {
            if (it < n_valid) {
                // ---- phase A: evaluate the filter, leave keep masks and counts in shared memory
                QE_RT_RESET();
                qe_mbar_wait(cb + O_FULL + 8u * st_a, par_a);
                QE_RT_ADD(9);
                const qe_u32 t = qe_lds_u32(cb + O_TILE + 4u * st_a);
                if (t >= p.ntiles) {
                    n_valid = it;
                } else {
                    const qe_u32 stage = dyn + st_a * (qe_u32)P::STAGE_BYTES;
                    const qe_i64 row0 = (qe_i64)t * TILE;
                    qe_u32 mk[STEPS];
                    if (row0 + TILE <= p.nrows) {
#pragma unroll
                        for (int s = 0; s < STEPS; s++) {
                            typename P::Row row;
                            typename P::Out o;
                            P::load_row(stage, wrow0, s * 32, row);
                            mk[s] = __ballot_sync(QE_FULL_MASK, P::eval(p, row, o));
                        }
                    } else {                                   // the last tile: rows past the end never pass
                        const qe_i64 live = p.nrows - row0 - wrow0 - (qe_i64)lane;
#pragma unroll
                        for (int s = 0; s < STEPS; s++) {
                            typename P::Row row;
                            typename P::Out o;
                            P::load_row(stage, wrow0, s * 32, row);
                            mk[s] = __ballot_sync(QE_FULL_MASK, P::eval(p, row, o) && (qe_i64)(s * 32) < live);
                        }
                    }
                    qe_u32 cnt = 0;
#pragma unroll
                    for (int s = 0; s < STEPS; s++) cnt += __popc(mk[s]);
                    if (LIST && cnt != 0u && cnt <= LIST_CAP) {
                        // few rows kept: their row numbers, packed in row order (steps that kept nothing cost one branch)
                        const qe_u32 la = cb + O_LIST + (qe_u32)LW * LIST_CAP * ((qe_u32)NCW * st_a + cw);
                        const qe_u32 lt = qe_lanemask_lt();
                        qe_u32 off = 0;
                        if (cnt > (qe_u32)STEPS) {
                            // more kept rows than steps: hardly a step is empty, the branches only cost (5-10 % kept)
#pragma unroll
                            for (int s = 0; s < STEPS; s++) {
                                qe_sts_entry_if<LW>((mk[s] >> lane) & 1u, la + (qe_u32)LW * (off + (qe_u32)__popc(mk[s] & lt)), (qe_u32)(s * 32) + lane);
                                off += (qe_u32)__popc(mk[s]);
                            }
                        } else {
#pragma unroll
                            for (int s = 0; s < STEPS; s++) {
                                if (mk[s] != 0u) {
                                    qe_sts_entry_if<LW>((mk[s] >> lane) & 1u, la + (qe_u32)LW * (off + (qe_u32)__popc(mk[s] & lt)), (qe_u32)(s * 32) + lane);
                                    off += (qe_u32)__popc(mk[s]);
                                }
                            }
                        }
                    }
                    const bool swept = NSW > 0 && cnt <= LIST_CAP;    // the sweepers store this slice's rows: this warp is done with the tile
                    __syncwarp();                                     // the list is written before lane 0 says so
                    if (lane == 0) {
                        if (!swept) {
                            const qe_u32 ma = cb + O_MASK + 4u * ((qe_u32)(TILE / 32) * st_a + cw * (qe_u32)STEPS);
                            if (STEPS % 4 == 0) {
#pragma unroll
                                for (int s = 0; s < STEPS; s += 4) qe_sts_u32x4(ma + 4u * s, mk[s], mk[s + 1], mk[s + 2], mk[s + 3]);
                            } else {
#pragma unroll
                                for (int s = 0; s < STEPS; s++) qe_sts_u32(ma + 4u * s, mk[s]);
                            }
                        }
                        qe_sts_u32(cb + O_CNT + 4u * ((qe_u32)NCW * st_a + cw), cnt);
                        qe_mbar_arrive(cb + O_EVAL + 8u * st_a);
                        if (swept) qe_mbar_arrive(cb + O_EMPTY + 8u * st_a);
                    }
                    __syncwarp();
                    QE_RT_ADD(10);
                }
                QE_RING_NEXT(st_a, par_a, 1u);
            }
            const int j = it - DEFER;
            if (j >= 0 && j < n_valid) {
                // ---- phase C: ordered scatter of tile j, whose prefix has had DEFER tile times to arrive
                QE_RT_RESET();
                const qe_u32 cnt_c = qe_lds_u32(cb + O_CNT + 4u * ((qe_u32)NCW * st_c + cw));      // this warp's own store of phase A
                const bool swept = NSW > 0 && cnt_c <= LIST_CAP;                   // nothing left to do here, the stage was released in phase A
                if (!swept && !(QE_DEBUG_FLAGS & 4)) qe_mbar_wait(cb + O_PREF + 8u * st_c, par_c);   // pref follows eval
                QE_RT_ADD(11);
                const qe_u32 cnt = (swept || (QE_DEBUG_FLAGS & 4)) ? 0u : cnt_c;   // 4: no phase C at all (timing only)
                if (cnt != 0u) {
                    const qe_u32 stage = dyn + st_c * (qe_u32)P::STAGE_BYTES;
                    const qe_u32 first_row = qe_lds_u32(cb + O_ROW0 + 4u * st_c) + (qe_u32)wrow0 + lane;
                    const qe_u64 wbase = qe_lds_u64(cb + O_WBASE + 8u * ((qe_u32)NCW * st_c + cw));   // output position of this warp's first kept row
                    const qe_u32 mk = cb + O_MASK + 4u * ((qe_u32)(TILE / 32) * st_c + cw * (qe_u32)STEPS);
                    const bool fits = qe_lds_u32(cb + O_FITS + 4u * st_c) != 0u;
                    if (LIST && fits && cnt <= LIST_CAP) {
                        const qe_u32 la = cb + O_LIST + (qe_u32)LW * LIST_CAP * ((qe_u32)NCW * st_c + cw);
                        if (p.sel) qe_ring_scatter_list<P, true>(p, stage, wrow0, first_row - lane, wbase, la, cnt);
                        else qe_ring_scatter_list<P, false>(p, stage, wrow0, first_row - lane, wbase, la, cnt);
                    } else if (fits) {
                        if (p.sel) qe_ring_scatter<P, STEPS, false, true>(p, stage, wrow0, first_row, wbase, mk, cnt);
                        else qe_ring_scatter<P, STEPS, false, false>(p, stage, wrow0, first_row, wbase, mk, cnt);
                    } else {
                        if (p.sel) qe_ring_scatter<P, STEPS, true, true>(p, stage, wrow0, first_row, wbase, mk, cnt);
                        else qe_ring_scatter<P, STEPS, true, false>(p, stage, wrow0, first_row, wbase, mk, cnt);
                    }
                }
                __syncwarp();
                QE_RT_ADD(12);
                if (lane == 0 && !swept) qe_mbar_arrive(cb + O_EMPTY + 8u * st_c);
                QE_RING_NEXT(st_c, par_c, 1u);
            }
            if (n_valid != 0x7fffffff && j + 1 >= n_valid) { if (cw == 0) QE_RT_FLUSH(9, 12); break; }
        }
    }
#undef QE_RING_NEXT
}

template <class P>
__device__ __forceinline__ void qe_project_body(const QeParams &p)
{
    constexpr int STEPS = P::STEPS, VEC = P::VEC;
    constexpr int WARP_ROWS = STEPS * 32 * VEC;
    constexpr int TILE = QE_WARPS * WARP_ROWS;
    const qe_u32 lane = qe_lane(), warp = qe_warp();

    for (qe_u32 tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x) {
        const qe_i64 warp_row0 = (qe_i64)tile * TILE + (qe_i64)warp * WARP_ROWS;
        typename P::Out outs[STEPS][VEC];
        qe_u32 keep[STEPS], inb[STEPS];
        qe_tile_eval<P>(p, warp_row0, outs, keep, inb);
#pragma unroll
        for (int s = 0; s < STEPS; s++) {
            const qe_i64 step_row0 = warp_row0 + (qe_i64)s * 32 * VEC;
            P::emit_dense(p, step_row0 + lane * VEC, outs[s], inb[s]);
#pragma unroll
            for (int b = 0; b < P::NBITCOLS; b++) {
                qe_u32 *bm = P::bitcol(p, b);
                if (bm) {
                    qe_u32 bits = 0;
#pragma unroll
                    for (int j = 0; j < VEC; j++) bits |= (P::bit(outs[s][j], b) ? 1u : 0u) << j;
                    qe_store_bits_dense<VEC>(bm, step_row0, bits & inb[s]);
                }
            }
        }
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) *p.out_count = (qe_u64)p.nrows;
}

// K2, ring variant: the columns of a tile arrive in shared-memory stages by bulk copy (one producer warp, static
// tile -> CTA assignment), NCW consumer warps evaluate one row per lane and step from shared memory and store the
// dense outputs (one coalesced 256-byte store per 8-byte output and warp step, validity / BOOLEAN words by ballot).
// The bytes in flight are bounded by shared memory instead of by the registers of four small CTAs.
template <class P>
__device__ __forceinline__ void qe_project_ring_body(const QeParams &p)
{
    constexpr int NST = P::NSTAGES, NCW = P::NCW, TILE = P::TILE_ROWS;
    constexpr int WROWS = TILE / NCW, STEPS = WROWS / 32;
    static_assert(WROWS % 32 == 0 && NCW >= 1 && NCW <= 31 && NST >= 2, "projection ring shape");

    extern __shared__ __align__(128) unsigned char qe_dyn_smem[];
    __shared__ __align__(8) qe_u64 s_bar[2 * NST];            // full[NST], empty[NST]
    __shared__ qe_u32 s_tile[NST];
    const qe_u32 warp = qe_warp(), lane = qe_lane();
    const qe_u32 bar = qe_opaque_u32(qe_smem_addr(s_bar));
    const qe_u32 dyn = qe_opaque_u32(qe_smem_addr(qe_dyn_smem));
    const qe_u32 tl = qe_opaque_u32(qe_smem_addr(s_tile));

    if (threadIdx.x == 0) {
#pragma unroll
        for (int st = 0; st < NST; st++) {
            qe_mbar_init(bar + 8u * st, 1);
            qe_mbar_init(bar + 8u * (NST + st), NCW);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    if (warp == 0) {
        if (lane == 0) {
            const qe_u64 policy = qe_policy_evict_first();
            qe_u32 st = 0, par = 0;
            for (qe_u32 it = 0;; it++) {
                if (it >= (qe_u32)NST) qe_mbar_wait(bar + 8u * (NST + st), par ^ 1u);
                const qe_u32 t = blockIdx.x + it * gridDim.x;
                qe_sts_u32(tl + 4u * st, t);
                const qe_u32 full = bar + 8u * st;
                if (t >= p.ntiles) {                          // end marker
                    qe_mbar_arrive(full);
                    break;
                }
                qe_mbar_expect_tx(full, (qe_u32)P::STAGE_TX);
                P::issue(p, (qe_i64)t * TILE, dyn + st * (qe_u32)P::STAGE_BYTES, full, policy);
                if (++st == (qe_u32)NST) { st = 0; par ^= 1u; }
            }
        }
    } else {
        const qe_u32 cw = warp - 1u;
        const int wrow0 = (int)cw * WROWS;
        qe_u32 st = 0, par = 0;
        for (;;) {
            qe_mbar_wait(bar + 8u * st, par);
            const qe_u32 t = qe_lds_u32(tl + 4u * st);
            if (t >= p.ntiles) break;
            const qe_u32 stage = dyn + st * (qe_u32)P::STAGE_BYTES;
            const qe_i64 row0 = (qe_i64)t * TILE + wrow0;     // first row of this warp's slice
#pragma unroll
            for (int s = 0; s < STEPS; s++) {
                typename P::Row row;
                typename P::Out o;
                P::load_row(stage, wrow0, s * 32, row);
                (void)P::eval(p, row, o);
                const qe_i64 r = row0 + s * 32 + (qe_i64)lane;
                const bool inb = r < p.nrows;
                if (inb) P::emit(p, (qe_u64)r, o);
#pragma unroll
                for (int b = 0; b < P::NBITCOLS; b++) {
                    qe_u32 *bm = P::bitcol(p, b);
                    const qe_u32 w = __ballot_sync(QE_FULL_MASK, inb && P::bit(o, b));
                    if (bm && lane == 0) bm[(row0 + s * 32) >> 5] = w;
                }
            }
            __syncwarp();
            if (lane == 0) qe_mbar_arrive(bar + 8u * (NST + st));
            if (++st == (qe_u32)NST) { st = 0; par ^= 1u; }
        }
    }
}

template <class P>
__device__ __forceinline__ void qe_agg_body(const QeParams &p)
{
    constexpr int STEPS = P::STEPS, VEC = P::VEC;
    constexpr int WARP_ROWS = STEPS * 32 * VEC;
    constexpr int TILE = QE_WARPS * WARP_ROWS;
    constexpr int TPS = QE_AGG_SUB_ROWS / TILE;               // tiles per sub-tile
    static_assert(QE_AGG_SUB_ROWS % TILE == 0 && TPS >= 1, "a sub-tile is a whole number of tiles");
    __shared__ QeAcc s_slot[QE_WARPS];
    const qe_u32 warp = qe_warp();

    qe_i64 kept = 0;
    const qe_u32 nsub = (p.ntiles + (qe_u32)TPS - 1u) / (qe_u32)TPS;
    // One partial per sub-tile of QE_AGG_SUB_ROWS rows, folded in an order that depends on the tile shape only: a
    // thread adds its rows of the sub-tile's tiles in ascending order, then the fixed lane tree and the warps 0..7
    // in sequence.  Which CTA takes which sub-tile does not matter (the partial is stored by sub-tile index), so
    // neither the grid size nor the partitioning of the table changes a SUM.
    for (qe_u32 sub = blockIdx.x; sub < nsub; sub += gridDim.x) {
        typename P::Accs accs;
        P::acc_init(accs);
        const qe_u32 t0 = sub * (qe_u32)TPS;
        const qe_u32 t1 = t0 + (qe_u32)TPS < p.ntiles ? t0 + (qe_u32)TPS : p.ntiles;
        for (qe_u32 tile = t0; tile < t1; tile++) {
            const qe_i64 warp_row0 = (qe_i64)tile * TILE + (qe_i64)warp * WARP_ROWS;
            typename P::Out outs[STEPS][VEC];
            qe_u32 keep[STEPS], inb[STEPS];
            qe_tile_eval<P>(p, warp_row0, outs, keep, inb);
#pragma unroll
            for (int s = 0; s < STEPS; s++) {
#pragma unroll
                for (int j = 0; j < VEC; j++) {
                    const bool k = (keep[s] >> j) & 1u;
                    P::acc_add(p, accs, outs[s][j], k);
                    kept += k ? 1 : 0;
                }
            }
        }
        P::acc_store(p, accs, s_slot, (qe_i64)sub);
    }
    // rows that reached the aggregation (diagnostic; the result row count is 1)
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) kept += __shfl_down_sync(QE_FULL_MASK, kept, o);
    if (qe_lane() == 0 && kept) atomicAdd((unsigned long long *)p.out_count, (unsigned long long)kept);
}

// ------------------------------------------------------------------ group-by
// GroupByAggregationOperator (M/operator/GroupByAggregationOperator.kt:21-75): a LinkedHashMap from the
// key columns of a row (boxed values compared with contentEquals: Double.equals, so NaN == NaN and
// -0.0 != 0.0; null is a key value like any other) to one Accumulator per aggregate; nulls are skipped
// (:38-43); rows come out in order of first appearance.
//
// Here: the groups live in an open-addressing table in HBM (g_state / g_keys / g_nulls), keyed by the
// canonical bits of the key values.  Accumulators are 8-byte words; word 0 holds the smallest row index of the
// group (first appearance) in its high half and the group's row count in its low half (a table has < 2^32
// rows), then come the words the aggregates need.  Two kernel bodies:
//   qe_groupby_private_body  few groups: every THREAD owns a private copy of the accumulator words of up
//                            to LMAX groups in shared memory (no atomics, no conflicts: lane-consecutive
//                            words), a shared-memory key cache maps key -> local group; a fixed-order
//                            reduce per CTA and a fixed-order merge over the CTAs follow, so the result
//                            does not depend on scheduling.
//   qe_groupby_atomic_body   any number of groups: global atomics on g_acc (SUM order not reproducible).
// What a group-by program P adds to load / eval:
//   static constexpr int NKEYS, NWORDS, LMAX, CSLOTS;
//   static void key_pack(const Out&, qe_u64 (&k)[NKEYS], qe_u32 &nulls);
//   static qe_u64 word_init(int w);  static int word_kind(int w);
//   static constexpr int GTHREADS;                                 private variant: threads per CTA (256 or 512)
//   static void acc_private(qe_u64 *a, const Out&, qe_u64 row);    word w of this thread's copy at a[w * LMAX * GTHREADS]
//   template <int AS> static void acc_atomic(qe_u64 *a, const Out&, qe_u64 row);     word w at a[w * AS]
#define QE_GK_ADD_F64 0
#define QE_GK_ADD_I64 1
#define QE_GK_MIN_F64 2
#define QE_GK_MAX_F64 3
#define QE_GK_MIN_U64 4
#define QE_GK_FIRST_ROWS 5     /* word 0: first row (u32, min) << 32 | rows of the group (u32, sum); a table has < 2^32 rows */

// Double.equals / hashCode compare doubleToLongBits: one NaN, and -0.0 differs from 0.0
__device__ __forceinline__ qe_u64 qe_key_f64(double x)
{
    return x != x ? 0x7ff8000000000000ULL : (qe_u64)qe_f64_bits(x);
}

__device__ __forceinline__ qe_u64 qe_group_combine(int kind, qe_u64 a, qe_u64 b)
{
    switch (kind) {
    case QE_GK_ADD_F64: return (qe_u64)qe_f64_bits(qe_add(qe_bits_f64((qe_i64)a), qe_bits_f64((qe_i64)b)));
    case QE_GK_ADD_I64: return a + b;
    case QE_GK_MIN_F64: return (qe_u64)qe_f64_bits(qe_math_min(qe_bits_f64((qe_i64)a), qe_bits_f64((qe_i64)b)));
    case QE_GK_MAX_F64: return (qe_u64)qe_f64_bits(qe_math_max(qe_bits_f64((qe_i64)a), qe_bits_f64((qe_i64)b)));
    case QE_GK_FIRST_ROWS: {
        const qe_u64 fa = a >> 32, fb = b >> 32;
        return ((fa < fb ? fa : fb) << 32) | ((a + b) & 0xffffffffULL);
    }
    default:            return a < b ? a : b;
    }
}

template <int NKEYS>
__device__ __forceinline__ qe_u32 qe_group_hash(const qe_u64 (&k)[NKEYS], qe_u32 nulls)
{
    // The high bits must come DOWN before the multiply: small-integer doubles differ only in their top ~12
    // bits, and a multiply alone never moves those into the low bits the table index is taken from
    // (measured: every key in one probe chain).
    qe_u64 h = 0x9e3779b97f4a7c15ULL ^ (qe_u64)nulls;
#pragma unroll
    for (int i = 0; i < NKEYS; i++) {
        h ^= k[i];
        h ^= h >> 32;
        h *= 0xff51afd7ed558ccdULL;
        h ^= h >> 29;
    }
    return (qe_u32)h;
}

// all 64 bits of the same hash (its low half is qe_group_hash): four 16-bit fields = four independent home slots
template <int NKEYS>
__device__ __forceinline__ qe_u64 qe_group_hash64(const qe_u64 (&k)[NKEYS], qe_u32 nulls)
{
    qe_u64 h = 0x9e3779b97f4a7c15ULL ^ (qe_u64)nulls;
#pragma unroll
    for (int i = 0; i < NKEYS; i++) {
        h ^= k[i];
        h ^= h >> 32;
        h *= 0xff51afd7ed558ccdULL;
        h ^= h >> 29;
    }
    return h;
}

__device__ __forceinline__ qe_u32 qe_ld_acquire_u32(const qe_u32 *p)
{
    qe_u32 v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

__device__ __forceinline__ void qe_st_release_u32(qe_u32 *p, qe_u32 v)
{
    asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// Slot of the key in the global table, inserting it if it is new.  -1: the table is full.
// Shape of the loop: no thread leaves it from inside the branch that holds a half-written slot -- every
// lane runs to the end of the iteration, so a lane that waits for a slot owned by another lane of its own
// warp cannot keep that lane from finishing (the usual intra-warp spin-lock hazard).  Spins are bounded:
// a run that hits the bound is discarded through the overflow flags instead of hanging the device.
template <int NKEYS, bool COUNT = true>
__device__ __noinline__ int qe_group_global_slot(const QeParams &p, const qe_u64 (&k)[NKEYS], qe_u32 nulls, qe_u32 h)
{
    qe_u32 s = h & p.g_slot_mask;
    qe_u32 probe = 0, spins = 0;
    int found = -2;
    while (found == -2) {
        const qe_u32 st = qe_ld_acquire_u32(p.g_state + s);
        if (st == 2u) {
            bool same = p.g_nulls[s] == nulls;
#pragma unroll
            for (int i = 0; i < NKEYS; i++) same = same && p.g_keys[(size_t)s * NKEYS + i] == k[i];
            if (same) {
                found = (int)s;
            } else {
                s = (s + 1u) & p.g_slot_mask;
                if (++probe > p.g_slot_mask) found = -1;
            }
        } else if (st == 0u) {
            if (atomicCAS(p.g_state + s, 0u, 1u) == 0u) {
                // (COUNT = false: the caller counts a whole partition's groups at once -- a million increments of one
                // word take milliseconds)
                if (COUNT && atomicAdd(p.g_ctl, 1u) >= p.g_max_groups) atomicOr(p.g_ctl + 1, 2u);   // over the limit: the run is discarded
#pragma unroll
                for (int i = 0; i < NKEYS; i++) p.g_keys[(size_t)s * NKEYS + i] = k[i];
                p.g_nulls[s] = nulls;
                qe_st_release_u32(p.g_state + s, 2u);
                found = (int)s;
            }
        } else if (++spins > (1u << 24)) {
            found = -1;                                       // the key of this slot never got published
        }
    }
    if (found < 0) atomicOr(p.g_ctl + 1, 2u);
    return found;
}

// The CTA's key cache of the private variant: key -> local group (index of the thread-private accumulators).
// Sized so that nearly every key sits in its first probe slot (CS >= 32 x LMAX): the per-row path is then
// straight-line code, the probing loop below runs for new and for displaced keys only.
template <int NKEYS, int CS, int LMAX>
struct QeGroupCache {
    qe_u64 keys[CS][NKEYS];
    qe_u32 nulls[CS];
    int local[CS];
    qe_u32 state[CS];           // 0 empty, 1 being written, 2 ready
    qe_u32 slot[LMAX];          // table slot of every local group
    qe_u32 nlocal, over;
};

template <int NKEYS, int CS, int LMAX>
__device__ __noinline__ int qe_group_local_slow(const QeParams &p, QeGroupCache<NKEYS, CS, LMAX> *gcp, const qe_u64 (&k)[NKEYS],
                                                qe_u32 nulls, qe_u32 h)
{
    volatile QeGroupCache<NKEYS, CS, LMAX> &gc = *gcp;
    int l = -1;
    qe_u32 c = h & (qe_u32)(CS - 1);
    int probe = 0;
    qe_u32 spins = 0;
    bool done = false;
    while (!done) {                                           // same loop shape as qe_group_global_slot
        const qe_u32 st = gc.state[c];
        if (gc.over) {
            done = true;                                      // the run is discarded anyway
        } else if (st == 2u) {
            bool same = gc.nulls[c] == nulls;
#pragma unroll
            for (int i = 0; i < NKEYS; i++) same = same && gc.keys[c][i] == k[i];
            if (same) {
                l = gc.local[c];
                done = true;
            } else {
                c = (c + 1u) & (qe_u32)(CS - 1);
                if (++probe >= CS) { gc.over = 1u; done = true; }
            }
        } else if (st == 0u) {
            if (atomicCAS(&gcp->state[c], 0u, 1u) == 0u) {
#pragma unroll
                for (int i = 0; i < NKEYS; i++) gc.keys[c][i] = k[i];
                gc.nulls[c] = nulls;
                int nl = (int)atomicAdd(&gcp->nlocal, 1u);
                if (nl >= LMAX) {
                    gc.over = 1u;
                    nl = -1;
                } else {
                    const int gs = qe_group_global_slot<NKEYS>(p, k, nulls, h);
                    gc.slot[nl] = gs < 0 ? 0xffffffffu : (qe_u32)gs;
                }
                gc.local[c] = nl;
                __threadfence_block();
                gc.state[c] = 2u;
                l = nl;
                done = true;
            }
        } else if (++spins > (1u << 24)) {
            gc.over = 1u;
            done = true;
        }
    }
    return l;
}

// Key cache of the shared-memory variants with copies (qe_groupby_copies_body): key -> local group.  These kernels are
// bound by the number of shared-memory wavefronts per row (ncu: l1tex data-pipe wavefronts 85-95 % of peak), and a miss
// in the straight-line lookup sends the whole warp through a loop, so the cache is built for two things: few loads per
// lookup and (nearly) no misses, whatever the keys look like.
//   * a slot is ONE 32-bit entry: ready | young | local group (13 bits) | null flags of the key (8) | 9 hash bits;
//     0 = empty, 1 = being written.  The key itself is stored once per GROUP (gkeys[local]).
//   * a bucket is two adjacent entries, read with one 64-bit load.  A key lives in the first free slot of its home
//     bucket, else of the following buckets.  With >= 8 slots per group a bucket holds 0.25 keys on average and fewer
//     than one key in a thousand sits outside its home bucket.
//   * lookup: load the bucket, pick the entry whose null flags and hash bits match, load that group's key and compare
//     (two dependent loads: 6 + 6 wavefronts per 32 rows at random addresses; round 1's cache took four loads, 16.5).
//     Everything else -- new keys, the rare displaced key, a 9-bit false positive next to the right entry -- takes
//     qe_group_local_slow2.
// YOUNG marks a group inserted in the current epoch (QE_GM_EPOCH_TILES tiles of the CTA between two barriers): only rows
// that see it, or come out of the slow path, update the group's first-row word -- in later epochs every row of the CTA
// has a larger row index than the row that inserted the group.
#define QE_GM_READY 0x80000000u
#define QE_GM_YOUNG 0x40000000u
#define QE_GM_MATCH 0x8001ffffu      /* ready + null flags + hash bits */
#define QE_GM_EPOCH_TILES 32u

template <int NKEYS, int CS, int LMAX>
struct QeGroupCache2 {
    qe_u32 entry[CS];           // CS / 2 buckets
    qe_u64 gkeys[LMAX][NKEYS];  // key of every local group
    qe_u32 slot[LMAX];          // table slot of every local group
    qe_u32 nlocal, over, inserted, pad;
};

__device__ __forceinline__ qe_u32 qe_group_entry_tag(qe_u64 h, qe_u32 nulls)
{
    return QE_GM_READY | (nulls << 9) | ((qe_u32)(h >> 40) & 0x1ffu);
}

// finds the key or inserts it; the local group, or -1 when the run has to be discarded (gc.over)
template <int NKEYS, int CS, int LMAX, bool GLOBAL = true>
__device__ __noinline__ int qe_group_local_slow2(const QeParams &p, QeGroupCache2<NKEYS, CS, LMAX> *gcp, const qe_u64 (&k)[NKEYS],
                                                 qe_u32 nulls, qe_u64 h)
{
    static_assert(LMAX <= 8192 && (CS & (CS - 1)) == 0 && CS >= 2, "13 bits of local group; buckets of two");
    volatile QeGroupCache2<NKEYS, CS, LMAX> &gc = *gcp;
    const qe_u32 tag = qe_group_entry_tag(h, nulls);
    int l = -1;
    qe_u32 c = ((qe_u32)h & (qe_u32)(CS / 2 - 1)) * 2u;      // first slot of the home bucket
    int probe = 0;
    qe_u32 spins = 0;
    bool done = false;
    while (!done) {                                           // same loop shape as qe_group_global_slot
        const qe_u32 e = gc.entry[c];
        if (gc.over) {
            done = true;                                      // the run is discarded anyway
        } else if (e & QE_GM_READY) {
            bool same = (e & QE_GM_MATCH) == tag;
            if (same) {
                const int cand = (int)((e >> 17) & 0x1fffu);
#pragma unroll
                for (int i = 0; i < NKEYS; i++) same = same && gc.gkeys[cand][i] == k[i];
                if (same) l = cand;
            }
            if (same) {
                done = true;
            } else {
                c = (c + 1u) & (qe_u32)(CS - 1);
                if (++probe >= CS) { gc.over = 1u; done = true; }
            }
        } else if (e == 0u) {
            if (atomicCAS(&gcp->entry[c], 0u, 1u) == 0u) {
                const int nl = (int)atomicAdd(&gcp->nlocal, 1u);
                if (nl >= LMAX) {
                    gc.over = 1u;                             // the slot stays "being written": whoever waits for it sees `over`
                } else {
#pragma unroll
                    for (int i = 0; i < NKEYS; i++) gc.gkeys[nl][i] = k[i];
                    if (GLOBAL) {                             // (the two-pass variant enters its groups into the table per partition)
                        const int gs = qe_group_global_slot<NKEYS>(p, k, nulls, (qe_u32)h);
                        gc.slot[nl] = gs < 0 ? 0xffffffffu : (qe_u32)gs;
                    }
                    gc.inserted = 1u;
                    __threadfence_block();
                    gc.entry[c] = tag | QE_GM_YOUNG | ((qe_u32)nl << 17);
                    l = nl;
                }
                done = true;
            }
        } else if (++spins > (1u << 24)) {
            gc.over = 1u;
            done = true;
        }
    }
    return l;
}

template <class P>
__device__ __forceinline__ void qe_groupby_private_body(const QeParams &p)
{
    constexpr int STEPS = P::STEPS, VEC = P::VEC, NKEYS = P::NKEYS, NWORDS = P::NWORDS, LMAX = P::LMAX, CS = P::CSLOTS;
    constexpr int GT = P::GTHREADS, GW = GT / 32;             // threads / warps per CTA: 512 / 16 when the groups fit, else 256 / 8
    constexpr int WARP_ROWS = STEPS * 32 * VEC;
    constexpr int TILE = GW * WARP_ROWS;
    extern __shared__ __align__(16) unsigned char qe_gsmem[];
    qe_u64 *s_acc = (qe_u64 *)qe_gsmem;                       // [NWORDS][LMAX][GT]
    __shared__ QeGroupCache<NKEYS, CS, LMAX> gc;
    const qe_u32 warp = qe_warp(), lane = qe_lane(), tid = threadIdx.x;

    for (int i = (int)tid; i < CS; i += GT) gc.state[i] = 0u;
    if (tid == 0) { gc.nlocal = 0u; gc.over = 0u; }
#pragma unroll 1
    for (int w = 0; w < NWORDS; w++) {
        const qe_u64 init = P::word_init(w);
#pragma unroll 1
        for (int l = 0; l < LMAX; l++) s_acc[((size_t)w * LMAX + l) * GT + tid] = init;
    }
    __syncthreads();

    // Only 8 warps fit beside the accumulators, so the loads of the NEXT tile are issued before this tile is
    // evaluated and accumulated (register double buffer): the memory latency hides behind the shared-memory
    // read-modify-write chains instead of adding to them.
    typename P::Row rows[STEPS][VEC];
    if (blockIdx.x < p.ntiles) {
#pragma unroll
        for (int s = 0; s < STEPS; s++)
            P::load(p, (qe_i64)blockIdx.x * TILE + (qe_i64)warp * WARP_ROWS + (qe_i64)s * 32 * VEC + lane * VEC, rows[s]);
    }
    for (qe_u32 tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x) {
        const qe_i64 warp_row0 = (qe_i64)tile * TILE + (qe_i64)warp * WARP_ROWS;
        typename P::Row next[STEPS][VEC];
        // the column allocations are padded by whole tiles, but not by a whole grid of them: clamp
        const qe_u32 ntile = tile + gridDim.x < p.ntiles ? tile + gridDim.x : tile;
#pragma unroll
        for (int s = 0; s < STEPS; s++)
            P::load(p, (qe_i64)ntile * TILE + (qe_i64)warp * WARP_ROWS + (qe_i64)s * 32 * VEC + lane * VEC, next[s]);
        typename P::Out outs[STEPS][VEC];
        qe_u32 keep[STEPS];
        qe_rows_eval<P>(p, warp_row0, rows, outs, keep);
        // Three passes over the thread's rows (as in the register variant): probe the key cache for every row with
        // unconditional reads, so that they are all in flight together; then the rare misses; then the updates.
        const volatile QeGroupCache<NKEYS, CS, LMAX> &vc = gc;
        int lg[STEPS][VEC];
        qe_u32 miss = 0;
#pragma unroll
        for (int s = 0; s < STEPS; s++) {
#pragma unroll
            for (int j = 0; j < VEC; j++) {
                const bool kept = (keep[s] >> j) & 1u;
                qe_u64 k[NKEYS];
                qe_u32 nulls;
                P::key_pack(outs[s][j], k, nulls);
                const qe_u32 c = qe_group_hash<NKEYS>(k, nulls) & (qe_u32)(CS - 1);
                // state first: a slot seen ready (2) never changes again, and a thread's shared-memory reads complete in order
                bool hit = vc.state[c] == 2u;
                hit &= vc.nulls[c] == nulls;
#pragma unroll
                for (int i = 0; i < NKEYS; i++) hit &= vc.keys[c][i] == k[i];
                const int lc = vc.local[c];
                lg[s][j] = (kept && hit) ? lc : -1;
                miss |= (kept && !hit) ? (1u << (s * VEC + j)) : 0u;
            }
        }
        if (miss) {
#pragma unroll
            for (int s = 0; s < STEPS; s++) {
#pragma unroll
                for (int j = 0; j < VEC; j++) {
                    if (!((miss >> (s * VEC + j)) & 1u)) continue;
                    qe_u64 k[NKEYS];
                    qe_u32 nulls;
                    P::key_pack(outs[s][j], k, nulls);
                    lg[s][j] = qe_group_local_slow<NKEYS, CS, LMAX>(p, &gc, k, nulls, qe_group_hash<NKEYS>(k, nulls));
                }
            }
        }
#pragma unroll
        for (int s = 0; s < STEPS; s++) {
#pragma unroll
            for (int j = 0; j < VEC; j++) {
                const int l = lg[s][j];
                if (l >= 0) {
                    const qe_u64 row = (qe_u64)p.sel_base + (qe_u64)(warp_row0 + (qe_i64)s * 32 * VEC + lane * VEC + j);
                    P::acc_private(s_acc + (size_t)l * GT + tid, outs[s][j], row);
                }
            }
        }
#pragma unroll
        for (int s = 0; s < STEPS; s++)
#pragma unroll
            for (int j = 0; j < VEC; j++) rows[s][j] = next[s][j];
    }
    __syncthreads();
    // ---- fixed-order reduce of the GT private copies: (group, word) pairs round-robin over the warps
    const int nlocal = (int)(gc.nlocal < (qe_u32)LMAX ? gc.nlocal : (qe_u32)LMAX);
    for (int idx = (int)warp; idx < nlocal * NWORDS; idx += GW) {
        const int l = idx / NWORDS, w = idx - l * NWORDS;
        const int kind = P::word_kind(w);
        const qe_u64 *src = s_acc + ((size_t)w * LMAX + l) * GT;
        qe_u64 v = src[lane];
#pragma unroll
        for (int q = 1; q < GW; q++) v = qe_group_combine(kind, v, src[lane + 32 * q]);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const qe_u64 b = (qe_u64)__shfl_down_sync(QE_FULL_MASK, (unsigned long long)v, o);
            v = qe_group_combine(kind, v, b);
        }
        if (lane == 0) p.g_part[((size_t)blockIdx.x * LMAX + l) * NWORDS + w] = v;
    }
    qe_u32 *ps = p.g_part_slot + (size_t)blockIdx.x * (1 + LMAX);
    if (tid == 0) {
        ps[0] = (qe_u32)nlocal;
        if (gc.over) atomicOr(p.g_ctl + 1, 1u);
    }
    if ((int)tid < nlocal) ps[1 + tid] = gc.slot[tid];
}

// Register variant: at most RMAX groups (2-16, by accumulator width), every consumer THREAD keeps the words of all
// of them in registers and a row updates them with predicated instructions (`if (local == g) ...`, unrolled over
// g): no shared-memory read-modify-write chain per row, which is what bounds the private variant (ncu: short
// scoreboard 5.0 stalled warps per issue, LSU wavefronts 59 % of peak, 47 % of DRAM peak on the Q1 shape).  With
// the accumulators out of shared memory, shared memory holds a ring of bulk-copy stages like the filter kernel's:
// one producer warp (cp.async.bulk per column run, static tile -> CTA assignment so that the summation order
// depends on the grid only), NCW consumer warps, full / empty mbarriers, no look-back (nothing is ordered).
// Ends like the private variant: registers to shared memory as [word][group][thread], fixed-order reduce, one
// partial per CTA for qe_groupby_merge_kernel.  More than RMAX groups -> overflow flag, the caller moves on to the
// private variant; the flag is raised as soon as it is seen and every producer stops issuing tiles.
template <class P>
__device__ __forceinline__ void qe_groupby_reg_body(const QeParams &p)
{
    constexpr int NST = P::NSTAGES, NCW = P::NCW, TILE = P::TILE_ROWS, NKEYS = P::NKEYS, NWORDS = P::NWORDS, RMAX = P::RMAX, CS = P::RCSLOTS;
    constexpr int WROWS = TILE / NCW, STEPS = WROWS / 32, GT = NCW * 32;
    static_assert(WROWS % 32 == 0 && NCW >= 1 && NCW <= 31 && NST >= 2, "register group-by shape");
    static_assert((size_t)NWORDS * RMAX * GT * 8 <= (size_t)NST * P::STAGE_BYTES, "the final reduce reuses the stages");

    extern __shared__ __align__(128) unsigned char qe_dyn_smem[];
    __shared__ __align__(8) qe_u64 s_bar[2 * NST];            // full[NST], empty[NST]
    __shared__ qe_u32 s_tile[NST];
    __shared__ QeGroupCache<NKEYS, CS, RMAX> gc;
    const qe_u32 warp = qe_warp(), lane = qe_lane(), tid = threadIdx.x;
    const qe_u32 bar = qe_opaque_u32(qe_smem_addr(s_bar));
    const qe_u32 dyn = qe_opaque_u32(qe_smem_addr(qe_dyn_smem));
    const qe_u32 tl = qe_opaque_u32(qe_smem_addr(s_tile));

    for (int i = (int)tid; i < CS; i += (int)blockDim.x) gc.state[i] = 0u;
    if (tid == 0) {
        gc.nlocal = 0u;
        gc.over = 0u;
#pragma unroll
        for (int st = 0; st < NST; st++) {
            qe_mbar_init(bar + 8u * st, 1);
            qe_mbar_init(bar + 8u * (NST + st), NCW);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    qe_u32 rfirst[RMAX], rrows[RMAX];
    qe_u64 racc[RMAX][NWORDS];
#pragma unroll
    for (int g = 0; g < RMAX; g++) {
        rfirst[g] = 0xffffffffu;
        rrows[g] = 0u;
#pragma unroll
        for (int w = 0; w < NWORDS; w++) racc[g][w] = P::word_init(w);
    }

    if (warp == 0) {
        // ------------------------------------------------------------ producer
        if (lane == 0) {
            const qe_u64 policy = qe_policy_evict_first();
            qe_u32 st = 0, par = 0;
            for (qe_u32 it = 0;; it++) {
                if (it >= (qe_u32)NST) qe_mbar_wait(bar + 8u * (NST + st), par ^ 1u);
                qe_u32 t = blockIdx.x + it * gridDim.x;
                if (t < p.ntiles && (qe_ld_relaxed_u32(p.g_ctl + 1) & 1u)) t = p.ntiles;   // some CTA overflowed: the run is discarded
                qe_sts_u32(tl + 4u * st, t);
                const qe_u32 full = bar + 8u * st;
                if (t >= p.ntiles) {                          // end marker
                    qe_mbar_arrive(full);
                    break;
                }
                qe_mbar_expect_tx(full, (qe_u32)P::STAGE_TX);
                P::issue(p, (qe_i64)t * TILE, dyn + st * (qe_u32)P::STAGE_BYTES, full, policy);
                if (++st == (qe_u32)NST) { st = 0; par ^= 1u; }
            }
        }
    } else {
        // ----------------------------------------------------------- consumers
        const qe_u32 cw = warp - 1u;
        const int wrow0 = (int)cw * WROWS;
        qe_u32 st = 0, par = 0;
        bool flagged = false;
        for (;;) {
            qe_mbar_wait(bar + 8u * st, par);
            const qe_u32 t = qe_lds_u32(tl + 4u * st);
            if (t >= p.ntiles) break;
            const qe_u32 stage = dyn + st * (qe_u32)P::STAGE_BYTES;
            const volatile QeGroupCache<NKEYS, CS, RMAX> &vc = gc;
            if (!vc.over) {
                const qe_i64 live = p.nrows - (qe_i64)t * TILE - wrow0 - (qe_i64)lane;   // rows of this lane's column that exist
                // Three passes over the warp's STEPS rows, each straight-line code, so that the shared-memory reads of
                // all steps are in flight together (three warps per scheduler cannot hide a per-row chain of
                // evaluate -> probe -> compare -> update): (1) evaluate and probe the key cache once, every read
                // unconditional, (2) the rare misses -- new or displaced keys -- take the probing loop, (3) update.
                typename P::Out o[STEPS];
                int l[STEPS];
                qe_u32 miss = 0;
#pragma unroll
                for (int s = 0; s < STEPS; s++) {
                    typename P::Row row;
                    P::load_row(stage, wrow0, s * 32, row);
                    const bool keep = P::eval(p, row, o[s]) && (qe_i64)(s * 32) < live;
                    qe_u64 k[NKEYS];
                    qe_u32 nulls;
                    P::key_pack(o[s], k, nulls);
                    const qe_u32 c = qe_group_hash<NKEYS>(k, nulls) & (qe_u32)(CS - 1);
                    // state first: a slot seen ready (2) never changes again, and a thread's shared-memory reads complete in order
                    bool hit = vc.state[c] == 2u;
                    hit &= vc.nulls[c] == nulls;
#pragma unroll
                    for (int i = 0; i < NKEYS; i++) hit &= vc.keys[c][i] == k[i];
                    const int lc = vc.local[c];
                    l[s] = (keep && hit) ? lc : -1;
                    miss |= (keep && !hit) ? (1u << s) : 0u;
                }
                if (miss) {
#pragma unroll
                    for (int s = 0; s < STEPS; s++) {
                        if (!((miss >> s) & 1u)) continue;
                        qe_u64 k[NKEYS];
                        qe_u32 nulls;
                        P::key_pack(o[s], k, nulls);
                        l[s] = qe_group_local_slow<NKEYS, CS, RMAX>(p, &gc, k, nulls, qe_group_hash<NKEYS>(k, nulls));
                    }
                }
#pragma unroll
                for (int s = 0; s < STEPS; s++) {
                    const qe_u32 rowid = p.sel_base + t * (qe_u32)TILE + (qe_u32)(wrow0 + s * 32) + lane;
                    // SUM words: one DFMA by a 0.0 / 1.0 mask per group and word (bit-identical to the add: v * 1.0
                    // is v, v * 0.0 is a zero of either sign and leaves a sum that started at +0.0 unchanged) --
                    // unless an input is Inf / NaN, whose product with 0.0 would poison the other groups: selects then
                    if (P::all_finite(o[s])) {
#pragma unroll
                        for (int g = 0; g < RMAX; g++) P::acc_regs_fma(rfirst[g], rrows[g], racc[g], o[s], rowid, l[s], g);
                    } else {
#pragma unroll
                        for (int g = 0; g < RMAX; g++) P::acc_regs(rfirst[g], rrows[g], racc[g], o[s], rowid, l[s], g);
                    }
                }
            } else if (!flagged && cw == 0 && lane == 0) {
                atomicOr(p.g_ctl + 1, 1u);                    // tell every producer now
                flagged = true;
            }
            __syncwarp();
            if (lane == 0) qe_mbar_arrive(bar + 8u * (NST + st));
            if (++st == (qe_u32)NST) { st = 0; par ^= 1u; }
        }
    }
    __syncthreads();                                          // every stage has been read: the stages become the reduce buffer
    qe_u64 *s_acc = (qe_u64 *)qe_dyn_smem;                    // [NWORDS][RMAX][GT]
    if (warp > 0) {
        const qe_u32 ct = tid - 32u;
#pragma unroll
        for (int g = 0; g < RMAX; g++) {
            s_acc[((size_t)0 * RMAX + g) * GT + ct] = ((qe_u64)rfirst[g] << 32) | (qe_u64)rrows[g];
#pragma unroll
            for (int w = 1; w < NWORDS; w++) s_acc[((size_t)w * RMAX + g) * GT + ct] = racc[g][w];
        }
    }
    __syncthreads();
    // ---- fixed-order reduce of the GT thread copies: (group, word) pairs round-robin over the consumer warps
    const int nlocal = (int)(gc.nlocal < (qe_u32)RMAX ? gc.nlocal : (qe_u32)RMAX);
    if (warp > 0) {
        for (int idx = (int)warp - 1; idx < nlocal * NWORDS; idx += NCW) {
            const int l = idx / NWORDS, w = idx - l * NWORDS;
            const int kind = P::word_kind(w);
            const qe_u64 *src = s_acc + ((size_t)w * RMAX + l) * GT;
            qe_u64 v = src[lane];
#pragma unroll
            for (int q = 1; q < NCW; q++) v = qe_group_combine(kind, v, src[lane + 32 * q]);
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const qe_u64 b = (qe_u64)__shfl_down_sync(QE_FULL_MASK, (unsigned long long)v, o);
                v = qe_group_combine(kind, v, b);
            }
            if (lane == 0) p.g_part[((size_t)blockIdx.x * RMAX + l) * NWORDS + w] = v;
        }
    }
    qe_u32 *ps = p.g_part_slot + (size_t)blockIdx.x * (1 + RMAX);
    if (tid == 0) {
        ps[0] = (qe_u32)nlocal;
        if (gc.over) atomicOr(p.g_ctl + 1, 1u);
    }
    if ((int)tid < nlocal) ps[1 + tid] = gc.slot[tid];
}

// Middle variant: more groups than the private copies hold, few enough for ONE copy of the accumulators per CTA
// in shared memory (SLMAX groups, hundreds to thousands).  Rows update it with shared-memory atomics (native
// 32-bit for the first-row / row-count word and the counts, compare-and-swap loops for the f64 words), which
// keeps the traffic off the L2 atomic units the global variant saturates.  The CTA's copy is flushed as a
// partial and merged in CTA order like the private variant's; within a CTA the order of the additions into a
// SUM is not reproducible.
template <class P>
__device__ __forceinline__ void qe_groupby_shared_body(const QeParams &p)
{
    constexpr int STEPS = P::STEPS, VEC = P::VEC, NKEYS = P::NKEYS, NWORDS = P::NWORDS, LMAX = P::SLMAX, CS = P::SCSLOTS;
    constexpr int GT = P::GTHREADS, GW = GT / 32;
    constexpr int WARP_ROWS = STEPS * 32 * VEC;
    constexpr int TILE = GW * WARP_ROWS;
    // COPIES = 8: one copy per (lane & 7) -- four lanes of a warp and their peers in all warps share it; more groups fit.
    // COPIES = 32 ("lane" kernel): one copy of the words per LANE index, shared by the lanes of that index in all
    // warps ([group][word][lane]: lane-consecutive, conflict-free).  The same atomics, but two threads meet on a
    // word only when they have the same lane AND the same group: the CAS loops of the f64 words hardly ever retry.
    // SCOPIES = 0 ("warp" kernel): one copy per WARP, laid out [warp][word][group], updated with plain loads and stores --
    // no atomics at all.  Lanes of a warp that meet the same group in one step find out by writing their lane number to
    // a per-warp tag byte of the group and reading it back: the lane whose number stuck updates, the others try again.
    // (match.any does the same in one instruction, but its latency grows with the number of distinct values: ~500
    // cycles per step with 30 distinct groups among 32 lanes, 7-9 ms per 6e8 rows.)
    constexpr bool WARPC = P::SCOPIES == 0;
    constexpr int COPIES = WARPC ? GW : P::SCOPIES;
    typedef QeGroupCache<NKEYS, CS, LMAX> Cache;
    extern __shared__ __align__(16) unsigned char qe_gsmem[];
    qe_u64 *s_acc = (qe_u64 *)qe_gsmem;                       // [LMAX][NWORDS][COPIES], or [COPIES][NWORDS][LMAX] (WARPC)
    Cache &gc = *(Cache *)(qe_gsmem + (size_t)LMAX * NWORDS * COPIES * 8);
    volatile unsigned char *s_tag = qe_gsmem + (size_t)LMAX * NWORDS * COPIES * 8 + sizeof(Cache);   // WARPC: [warp][LMAX] lane numbers
    const qe_u32 warp = qe_warp(), lane = qe_lane(), tid = threadIdx.x;

    for (int i = (int)tid; i < CS; i += GT) gc.state[i] = 0u;
    if (tid == 0) { gc.nlocal = 0u; gc.over = 0u; }
    for (int i = (int)tid; i < LMAX * NWORDS * COPIES; i += GT) s_acc[i] = P::word_init(WARPC ? (i / LMAX) % NWORDS : (i / COPIES) % NWORDS);
    __syncthreads();

    typename P::Row rows[STEPS][VEC];
    if (blockIdx.x < p.ntiles) {
#pragma unroll
        for (int s = 0; s < STEPS; s++)
            P::load(p, (qe_i64)blockIdx.x * TILE + (qe_i64)warp * WARP_ROWS + (qe_i64)s * 32 * VEC + lane * VEC, rows[s]);
    }
    qe_u32 over_poll = 0;
    for (qe_u32 tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x) {
        // more groups than this variant holds, here or in another CTA: the run is discarded, stop reading.  Looked at
        // every 32nd tile only: the global flag is an L2 round trip in front of a tile of a few rows per thread (checked
        // every tile it cost the lane-copies kernel 20 %: 2.2 -> 2.7 ms).
        if ((++over_poll & 31u) == 0u &&
            (((const volatile qe_u32 &)gc.over) != 0u || (qe_ld_relaxed_u32(p.g_ctl + 1) & 1u))) {
            if (((const volatile qe_u32 &)gc.over) != 0u && lane == 0) atomicOr(p.g_ctl + 1, 1u);
            break;
        }
        const qe_i64 warp_row0 = (qe_i64)tile * TILE + (qe_i64)warp * WARP_ROWS;
        typename P::Row next[STEPS][VEC];
        const qe_u32 ntile = tile + gridDim.x < p.ntiles ? tile + gridDim.x : tile;
#pragma unroll
        for (int s = 0; s < STEPS; s++)
            P::load(p, (qe_i64)ntile * TILE + (qe_i64)warp * WARP_ROWS + (qe_i64)s * 32 * VEC + lane * VEC, next[s]);
        typename P::Out outs[STEPS][VEC];
        qe_u32 keep[STEPS];
        qe_rows_eval<P>(p, warp_row0, rows, outs, keep);
        // probes of all rows first (unconditional reads, all in flight together), misses second, updates third
        const volatile Cache &vc = gc;
        int lg[STEPS][VEC];
        qe_u32 miss = 0;
#pragma unroll
        for (int s = 0; s < STEPS; s++) {
#pragma unroll
            for (int j = 0; j < VEC; j++) {
                const bool kept = (keep[s] >> j) & 1u;
                qe_u64 k[NKEYS];
                qe_u32 nulls;
                P::key_pack(outs[s][j], k, nulls);
                const qe_u32 c = qe_group_hash<NKEYS>(k, nulls) & (qe_u32)(CS - 1);
                bool hit = vc.state[c] == 2u;
                hit &= vc.nulls[c] == nulls;
#pragma unroll
                for (int i = 0; i < NKEYS; i++) hit &= vc.keys[c][i] == k[i];
                const int lc = vc.local[c];
                lg[s][j] = (kept && hit) ? lc : -1;
                miss |= (kept && !hit) ? (1u << (s * VEC + j)) : 0u;
            }
        }
        if (miss) {
#pragma unroll
            for (int s = 0; s < STEPS; s++) {
#pragma unroll
                for (int j = 0; j < VEC; j++) {
                    if (!((miss >> (s * VEC + j)) & 1u)) continue;
                    qe_u64 k[NKEYS];
                    qe_u32 nulls;
                    P::key_pack(outs[s][j], k, nulls);
                    lg[s][j] = qe_group_local_slow<NKEYS, CS, LMAX>(p, &gc, k, nulls, qe_group_hash<NKEYS>(k, nulls));
                }
            }
        }
#pragma unroll
        for (int s = 0; s < STEPS; s++) {
#pragma unroll
            for (int j = 0; j < VEC; j++) {
                const int l = lg[s][j];
                const qe_u64 row = (qe_u64)p.sel_base + (qe_u64)(warp_row0 + (qe_i64)s * 32 * VEC + lane * VEC + j);
                if (WARPC) {
                    // lanes with the same group take turns: whose lane number sticks in the group's tag byte goes first
                    bool todo = l >= 0;
                    volatile unsigned char *tag = s_tag + (size_t)warp * LMAX + (l >= 0 ? l : 0);
                    while (__any_sync(QE_FULL_MASK, todo)) {
                        if (todo) *tag = (unsigned char)lane;
                        __syncwarp();
                        if (todo && *tag == (unsigned char)lane) {
                            P::template acc_plain<LMAX>(s_acc + (size_t)warp * NWORDS * LMAX + l, outs[s][j], row);
                            todo = false;
                        }
                        __syncwarp();
                    }
                } else if (l >= 0) {
                    P::template acc_atomic<COPIES>(s_acc + (size_t)l * NWORDS * COPIES + (lane & (qe_u32)(COPIES - 1)), outs[s][j], row);
                }
            }
        }
#pragma unroll
        for (int s = 0; s < STEPS; s++)
#pragma unroll
            for (int j = 0; j < VEC; j++) rows[s][j] = next[s][j];
    }
    __syncthreads();
    const int nlocal = (int)(gc.nlocal < (qe_u32)LMAX ? gc.nlocal : (qe_u32)LMAX);
    for (int i = (int)tid; i < nlocal * NWORDS; i += GT) {
        const int kind = P::word_kind(i % NWORDS);
        qe_u64 v;
        if (WARPC) {                                          // the warps' copies in warp order
            const size_t at = (size_t)(i % NWORDS) * LMAX + (size_t)(i / NWORDS);
            v = s_acc[at];
            for (int c = 1; c < COPIES; c++) v = qe_group_combine(kind, v, s_acc[(size_t)c * NWORDS * LMAX + at]);
        } else {
            v = s_acc[(size_t)i * COPIES];
            for (int c = 1; c < COPIES; c++) v = qe_group_combine(kind, v, s_acc[(size_t)i * COPIES + c]);   // the lane copies in lane order
        }
        p.g_part[(size_t)blockIdx.x * LMAX * NWORDS + i] = v;
    }
    qe_u32 *ps = p.g_part_slot + (size_t)blockIdx.x * (1 + LMAX);
    if (tid == 0) {
        ps[0] = (qe_u32)nlocal;
        if (gc.over) atomicOr(p.g_ctl + 1, 1u);
    }
    for (int i = (int)tid; i < nlocal; i += GT) ps[1 + i] = gc.slot[i];
}

// The shared-memory variants with COPIES = 32, 16, 8, 4, 2 or 1 copies of the accumulator words per CTA
// ([group][word][copy], copy = lane & (COPIES - 1): with 32 -- 16 for the 8-byte words -- every lane has its own bank),
// updated with shared-memory atomics like qe_groupby_shared_body, around the two-load key cache above.  What the
// kernel is bound by is the number of shared-memory wavefronts per row: a probe was 4 loads at random addresses (16.5
// wavefronts per 32 rows), now 2 (9.5); the look at the first-row word (3.5) is gone for all but young groups; more
// copies in the same shared memory (a slot is 12 bytes instead of 20, four slots per group instead of sixteen) cut the
// bank conflicts of the atomics.
template <class P>
__device__ __forceinline__ void qe_groupby_copies_body(const QeParams &p)
{
    constexpr int STEPS = P::STEPS, VEC = P::VEC, NKEYS = P::NKEYS, NWORDS = P::NWORDS, LMAX = P::SLMAX, CS = P::SCSLOTS;
    constexpr int GT = P::GTHREADS, GW = GT / 32;
    constexpr int WARP_ROWS = STEPS * 32 * VEC;
    constexpr int TILE = GW * WARP_ROWS;
    constexpr int COPIES = P::SCOPIES;
    static_assert(COPIES >= 1 && COPIES <= 32 && (COPIES & (COPIES - 1)) == 0, "copies per CTA: a power of two up to 32");
    static_assert(STEPS * VEC <= 32, "one miss bit per row of a thread");
    typedef QeGroupCache2<NKEYS, CS, LMAX> Cache;
    extern __shared__ __align__(16) unsigned char qe_gsmem[];
    qe_u64 *s_acc = (qe_u64 *)qe_gsmem;                       // [LMAX][NWORDS][COPIES]
    Cache &gc = *(Cache *)(qe_gsmem + (size_t)LMAX * NWORDS * COPIES * 8);
    const qe_u32 warp = qe_warp(), lane = qe_lane(), tid = threadIdx.x;

    for (int i = (int)tid; i < CS; i += GT) gc.entry[i] = 0u;
    if (tid == 0) { gc.nlocal = 0u; gc.over = 0u; gc.inserted = 0u; }
    for (int i = (int)tid; i < LMAX * NWORDS * COPIES; i += GT) s_acc[i] = P::word_init((i / COPIES) % NWORDS);
    __syncthreads();

    typename P::Row rows[STEPS][VEC];
    if (blockIdx.x < p.ntiles) {
#pragma unroll
        for (int s = 0; s < STEPS; s++)
            P::load(p, (qe_i64)blockIdx.x * TILE + (qe_i64)warp * WARP_ROWS + (qe_i64)s * 32 * VEC + lane * VEC, rows[s]);
    }
    qe_u32 it = 0;
    for (qe_u32 tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x) {
        // End of an epoch (every warp of the CTA runs the same tiles, so the barriers match).  More groups than this
        // variant holds, here or in another CTA: the run is discarded, stop reading.  Groups inserted during the
        // epoch stop being young: no warp is inside a tile now, and every later row of this CTA has a larger index.
        if ((++it & (QE_GM_EPOCH_TILES - 1u)) == 0u) {
            const bool mine = ((const volatile qe_u32 &)gc.over) != 0u;
            if (__syncthreads_or((mine || (tid == 0 && (qe_ld_relaxed_u32(p.g_ctl + 1) & 1u))) ? 1 : 0)) break;
            if (__syncthreads_or(((const volatile qe_u32 &)gc.inserted) != 0u ? 1 : 0)) {
                for (int i = (int)tid; i < CS; i += GT) {
                    const qe_u32 m = gc.entry[i];
                    if (m & QE_GM_YOUNG) gc.entry[i] = m & ~QE_GM_YOUNG;
                }
                if (tid == 0) gc.inserted = 0u;
                __syncthreads();
            }
        }
        const qe_i64 warp_row0 = (qe_i64)tile * TILE + (qe_i64)warp * WARP_ROWS;
        typename P::Row next[STEPS][VEC];
        const qe_u32 ntile = tile + gridDim.x < p.ntiles ? tile + gridDim.x : tile;
#pragma unroll
        for (int s = 0; s < STEPS; s++)
            P::load(p, (qe_i64)ntile * TILE + (qe_i64)warp * WARP_ROWS + (qe_i64)s * 32 * VEC + lane * VEC, next[s]);
        typename P::Out outs[STEPS][VEC];
        qe_u32 keep[STEPS];
        qe_rows_eval<P>(p, warp_row0, rows, outs, keep);
        // home buckets of all rows first (unconditional reads, all in flight together), the rare misses second, updates third
        const volatile Cache &vc = gc;
        int lg[STEPS][VEC];                                   // local group | young << 22, or -1
        qe_u32 miss = 0;
#pragma unroll
        for (int s = 0; s < STEPS; s++) {
#pragma unroll
            for (int j = 0; j < VEC; j++) {
                const bool kept = (keep[s] >> j) & 1u;
                qe_u64 k[NKEYS];
                qe_u32 nulls;
                P::key_pack(outs[s][j], k, nulls);
                const qe_u64 h = qe_group_hash64<NKEYS>(k, nulls);
                const qe_u32 tag = qe_group_entry_tag(h, nulls);
                const qe_u64 e = *(const volatile qe_u64 *)&gc.entry[((qe_u32)h & (qe_u32)(CS / 2 - 1)) * 2u];    // the home bucket
                const qe_u32 e0 = (qe_u32)e, e1 = (qe_u32)(e >> 32);
                const bool m0 = (e0 & QE_GM_MATCH) == tag, m1 = (e1 & QE_GM_MATCH) == tag;
                const qe_u32 pick = m0 ? e0 : e1;
                const int cand = (m0 || m1) ? (int)((pick >> 17) & 0x1fffu) : 0;
                bool hit = m0 || m1;
#pragma unroll
                for (int i = 0; i < NKEYS; i++) hit &= vc.gkeys[cand][i] == k[i];
                lg[s][j] = (kept && hit) ? (cand | (int)((pick >> 8) & 0x400000u)) : -1;
                miss |= (kept && !hit) ? (1u << (s * VEC + j)) : 0u;
            }
        }
        if (miss) {
            // A key that found its home bucket full sits in one of the next buckets (or, after a false match of the 9
            // hash bits, in the other slot of its home bucket): look there in line -- a single displaced key among a
            // few hundred sends a row of nearly every other tile this way, and the loop below costs ~10 x as much.
            // Measured alternatives (profiles/r02_groupby_copies_steps.jsonl): the next bucket loaded in the lookup above
            // for the rows that need it, predicated: +0.5 ms for every key set (the rows that do not need it wait for the
            // load as well); behind one vote per tile: +0.2 ms for every key set, none for displaced keys; here: nothing
            // for key sets without displaced keys, two shared-memory round trips for a displaced row's warp.
#pragma unroll
            for (int s = 0; s < STEPS; s++) {
#pragma unroll
                for (int j = 0; j < VEC; j++) {
                    if (!((miss >> (s * VEC + j)) & 1u)) continue;
                    qe_u64 k[NKEYS];
                    qe_u32 nulls;
                    P::key_pack(outs[s][j], k, nulls);
                    const qe_u64 h = qe_group_hash64<NKEYS>(k, nulls);
                    const qe_u32 tag = qe_group_entry_tag(h, nulls);
                    int l = -1;
                    {                                                          // the next bucket: one load for its two slots
                        const qe_u64 f = *(const volatile qe_u64 *)&gc.entry[((((qe_u32)h & (qe_u32)(CS / 2 - 1)) + 1u) & (qe_u32)(CS / 2 - 1)) * 2u];
                        const qe_u32 f0 = (qe_u32)f, f1 = (qe_u32)(f >> 32);
                        const bool n0 = (f0 & QE_GM_MATCH) == tag, n1 = (f1 & QE_GM_MATCH) == tag;
                        if (n0 || n1) {
                            const qe_u32 pick = n0 ? f0 : f1;
                            const int cand = (int)((pick >> 17) & 0x1fffu);
                            bool same = true;
#pragma unroll
                            for (int i = 0; i < NKEYS; i++) same &= vc.gkeys[cand][i] == k[i];
                            if (same) l = cand | (int)((pick >> 8) & 0x400000u);
                        }
                    }
                    for (qe_u32 d = 0; d < 8u && l < 0; d++) {                 // slots 0..7 from the home bucket on
                        const qe_u32 e = vc.entry[(((qe_u32)h & (qe_u32)(CS / 2 - 1)) * 2u + d) & (qe_u32)(CS - 1)];
                        if ((e & QE_GM_MATCH) != tag) continue;
                        const int cand = (int)((e >> 17) & 0x1fffu);
                        bool same = true;
#pragma unroll
                        for (int i = 0; i < NKEYS; i++) same &= vc.gkeys[cand][i] == k[i];
                        if (same) l = cand | (int)((e >> 8) & 0x400000u);
                    }
                    if (l < 0) {
                        l = qe_group_local_slow2<NKEYS, CS, LMAX>(p, &gc, k, nulls, h);
                        if (l >= 0) l |= 1 << 22;                             // out of the loop: treat as young
                    }
                    lg[s][j] = l;
                }
            }
        }
#pragma unroll
        for (int s = 0; s < STEPS; s++) {
#pragma unroll
            for (int j = 0; j < VEC; j++) {
                const int l = lg[s][j];
                if (l >= 0) {
                    const qe_u64 row = (qe_u64)p.sel_base + (qe_u64)(warp_row0 + (qe_i64)s * 32 * VEC + lane * VEC + j);
                    qe_u64 *a = s_acc + (size_t)(l & 0x3fffff) * NWORDS * COPIES + (lane & (qe_u32)(COPIES - 1));
                    if (l >> 22) atomicMin(((unsigned int *)a) + 1, (unsigned int)row);
                    // Row after row: overlapping the compare-and-swap loops of a thread's rows (all loads, then all CAS)
                    // widens the window between a load and its CAS and most of them fail -- 2.1 -> 2.9 ms at 256 groups,
                    // 6 ms at 32 (profiles/r02_groupby_copies_steps.jsonl, step 4).
                    P::template acc_atomic<COPIES, false>(a, outs[s][j], row);
                }
            }
        }
#pragma unroll
        for (int s = 0; s < STEPS; s++)
#pragma unroll
            for (int j = 0; j < VEC; j++) rows[s][j] = next[s][j];
    }
    __syncthreads();
    const int nlocal = (int)(gc.nlocal < (qe_u32)LMAX ? gc.nlocal : (qe_u32)LMAX);
    for (int i = (int)tid; i < nlocal * NWORDS; i += GT) {
        const int kind = P::word_kind(i % NWORDS);
        qe_u64 v = s_acc[(size_t)i * COPIES];
        for (int c = 1; c < COPIES; c++) v = qe_group_combine(kind, v, s_acc[(size_t)i * COPIES + c]);   // the copies in lane order
        p.g_part[(size_t)blockIdx.x * LMAX * NWORDS + i] = v;
    }
    qe_u32 *ps = p.g_part_slot + (size_t)blockIdx.x * (1 + LMAX);
    if (tid == 0) {
        ps[0] = (qe_u32)nlocal;
        if (gc.over) atomicOr(p.g_ctl + 1, 1u);
    }
    for (int i = (int)tid; i < nlocal; i += GT) ps[1 + i] = gc.slot[i];
}

// ---- more groups than a CTA's shared memory holds: two passes ------------------------------------------------------
// With global atomics every row costs ~6 L2 transactions at random addresses (probe of the table, first-row word, count,
// sum): 22-29 ms per 6e8 rows, whatever the group count.  Instead:
//   pass 1 (qe_groupby_scatter_body)  evaluates filter, keys and aggregate inputs once and writes each row as a record
//          (P::record_pack: row, null flags, key words, distinct input values) into one of NPART partitions chosen by
//          bits 49.. of the key hash.  Records are staged per partition in shared memory ([NPART][RW][PCAP] words) and
//          written out PCAP/2..PCAP at a time (one global atomic per flush reserves the space; word arrays are separate,
//          so pass 2 reads coalesced).  Order inside a partition is arbitrary -- the record carries its row.
//   pass 2 (qe_groupby_gather_body)   one CTA per partition: the copies variant's shared-memory table over the
//          partition's records (a partition holds 1/NPART of the groups), then the partition's groups -- complete, no
//          other CTA has any of their rows -- go into the table in HBM with plain stores.
// A partition that overflows its space (one key with a large share of the rows) or its CTA's table raises the overflow
// flag: the run is discarded and the caller goes on to the atomic variant.
#define QE_PT_VALID (1ull << 63)          /* word 0 of every record */

template <class P>
__device__ __forceinline__ void qe_groupby_scatter_flush(const QeParams &p, qe_u64 *st, volatile qe_u32 *fill, qe_u32 threshold)
{
    constexpr int NP = P::NPART, CAP = P::PCAP, RW = P::RW, GW = P::GTHREADS / 32;
    constexpr int WPI = 32 / CAP;                             // record words a warp copies per instruction (CAP lanes each)
    static_assert(CAP == 8 || CAP == 16 || CAP == 32, "staging records per partition");
    const qe_u32 warp = qe_warp(), lane = qe_lane();
    const size_t total = (size_t)NP * p.pt_cap;
    for (qe_u32 base = warp * 32u; base < (qe_u32)NP; base += (qe_u32)GW * 32u) {
        // Every lane reserves the space of its own partition: up to 32 global atomics in flight, one round trip (one
        // partition after the other it was 16 round trips per warp and tile: 40 ms per 6e8 rows instead of 8.4).
        // Measured and rejected: blocks of 128 records per CTA and partition, reserved one block ahead so that no tile
        // waits for L2 -- 148 x 512 write frontiers instead of 512, 13.3 ms.
        const qe_u32 f = fill[base + lane];
        const qe_u32 n = f < (qe_u32)CAP ? f : (qe_u32)CAP;      // rows that found the staging full wait for the next round
        const bool want = f >= threshold && n > 0u;
        qe_u32 g0 = 0;
        if (want && !(QE_DEBUG_FLAGS & 2048)) g0 = atomicAdd(p.pt_cursor + base + lane, n);
        const bool ok = want && g0 + n <= p.pt_cap;
        if (want && !ok) atomicOr(p.g_ctl + 1, 1u);
        qe_u32 m = __ballot_sync(QE_FULL_MASK, ok && !(QE_DEBUG_FLAGS & 2048));
        while (m) {
            const qe_u32 bit = (qe_u32)__ffs((int)m) - 1u;
            m &= m - 1u;
            const qe_u32 q = base + bit;
            const qe_u32 nn = __shfl_sync(QE_FULL_MASK, n, (int)bit), gg = __shfl_sync(QE_FULL_MASK, g0, (int)bit);
            const qe_u32 j = lane & (qe_u32)(CAP - 1);
#pragma unroll
            for (int w0 = 0; w0 < RW; w0 += WPI) {
                const int w = w0 + (int)(lane / (qe_u32)CAP);
                if (w < RW && j < nn) p.pt_words[(size_t)w * total + (size_t)q * p.pt_cap + gg + j] = st[((size_t)q * RW + w) * CAP + j];
            }
        }
        __syncwarp();
        if (want) fill[base + lane] = 0u;
    }
}

template <class P>
__device__ __forceinline__ void qe_groupby_scatter_body(const QeParams &p)
{
    constexpr int STEPS = P::STEPS, VEC = P::VEC, NKEYS = P::NKEYS, NP = P::NPART, CAP = P::PCAP, RW = P::RW;
    constexpr int GT = P::GTHREADS, GW = GT / 32;
    constexpr int WARP_ROWS = STEPS * 32 * VEC;
    constexpr int TILE = GW * WARP_ROWS;
    static_assert((NP & (NP - 1)) == 0 && NP % 32 == 0 && NP <= 512, "partitions: a power of two, 9 hash bits at most");
    static_assert(STEPS * VEC <= 32, "one pending bit per row of a thread");
    extern __shared__ __align__(16) unsigned char qe_gsmem[];
    qe_u64 *st = (qe_u64 *)qe_gsmem;                          // [NP][RW][CAP]
    volatile qe_u32 *fill = (volatile qe_u32 *)(st + (size_t)NP * RW * CAP);
    const qe_u32 warp = qe_warp(), lane = qe_lane(), tid = threadIdx.x;
    for (int i = (int)tid; i < NP; i += GT) fill[i] = 0u;
    __syncthreads();

    typename P::Row rows[STEPS][VEC];
    if (blockIdx.x < p.ntiles) {
#pragma unroll
        for (int s = 0; s < STEPS; s++)
            P::load(p, (qe_i64)blockIdx.x * TILE + (qe_i64)warp * WARP_ROWS + (qe_i64)s * 32 * VEC + lane * VEC, rows[s]);
    }
    qe_u32 it = 0;
    for (qe_u32 tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x) {
        if ((++it & 31u) == 0u && __syncthreads_or((tid == 0 && (qe_ld_relaxed_u32(p.g_ctl + 1) & 1u)) ? 1 : 0)) break;   // discarded
        const qe_i64 warp_row0 = (qe_i64)tile * TILE + (qe_i64)warp * WARP_ROWS;
        typename P::Row next[STEPS][VEC];
        const qe_u32 ntile = tile + gridDim.x < p.ntiles ? tile + gridDim.x : tile;
#pragma unroll
        for (int s = 0; s < STEPS; s++)
            P::load(p, (qe_i64)ntile * TILE + (qe_i64)warp * WARP_ROWS + (qe_i64)s * 32 * VEC + lane * VEC, next[s]);
        typename P::Out outs[STEPS][VEC];
        qe_u32 keep[STEPS];
        qe_rows_eval<P>(p, warp_row0, rows, outs, keep);
        qe_u64 rec[STEPS * VEC][RW];
        qe_u32 part[STEPS * VEC];
        qe_u32 pending = 0;
#pragma unroll
        for (int s = 0; s < STEPS; s++) {
#pragma unroll
            for (int j = 0; j < VEC; j++) {
                qe_u64 k[NKEYS];
                qe_u32 nulls;
                P::key_pack(outs[s][j], k, nulls);
                part[s * VEC + j] = (qe_u32)(qe_group_hash64<NKEYS>(k, nulls) >> 49) & (qe_u32)(NP - 1);
                P::record_pack(outs[s][j], p.sel_base + (qe_u32)(warp_row0 + (qe_i64)s * 32 * VEC + lane * VEC + j), k, nulls, rec[s * VEC + j]);
                rec[s * VEC + j][0] |= QE_PT_VALID;
                pending |= (((keep[s] >> j) & 1u) && !(QE_DEBUG_FLAGS & 1024)) ? (1u << (s * VEC + j)) : 0u;
            }
        }
        if ((QE_DEBUG_FLAGS & 1024) && rec[0][0] == 0x1234567ull) p.pt_words[tid] = rec[0][1];      // keeps the evaluation alive
        int more;
        do {
#pragma unroll
            for (int r = 0; r < STEPS * VEC; r++) {
                if ((pending >> r) & 1u) {
                    const qe_u32 pos = atomicAdd((qe_u32 *)fill + part[r], 1u);
                    if (pos < (qe_u32)CAP) {
#pragma unroll
                        for (int w = 0; w < RW; w++) st[((size_t)part[r] * RW + w) * CAP + pos] = rec[r][w];
                        pending &= ~(1u << r);
                    }
                }
            }
            __syncthreads();
            qe_groupby_scatter_flush<P>(p, st, fill, (qe_u32)CAP / 2u);
            more = __syncthreads_or(pending != 0u ? 1 : 0);
        } while (more);
#pragma unroll
        for (int s = 0; s < STEPS; s++)
#pragma unroll
            for (int j = 0; j < VEC; j++) rows[s][j] = next[s][j];
    }
    __syncthreads();
    qe_groupby_scatter_flush<P>(p, st, fill, 1u);
}

// two adjacent records' word w (16 bytes), streamed: no L1 allocation, fixed place in the instruction stream
__device__ __forceinline__ void qe_gather_load2(const qe_u64 *src, qe_u64 &a, qe_u64 &b)
{
    qe_i64 x, y;
    qe_ldg_i64x2((const qe_i64 *)src, x, y);
    a = (qe_u64)x;
    b = (qe_u64)y;
}

template <class P>
__device__ __forceinline__ void qe_groupby_gather_body(const QeParams &p)
{
    constexpr int NKEYS = P::NKEYS, NWORDS = P::NWORDS, LMAX = P::SLMAX, CS = P::SCSLOTS, NP = P::NPART, RW = P::RW;
    constexpr int GT = P::GTHREADS;
    constexpr int COPIES = P::SCOPIES;
    constexpr int PAIRS = RW <= 4 ? 2 : 1, R = 2 * PAIRS;     // records per thread and step, loaded as pairs of neighbours
    constexpr bool PREF = RW <= 6;                            // the next step's records live in registers too
    static_assert(COPIES >= 1 && COPIES <= 32 && (COPIES & (COPIES - 1)) == 0, "copies per CTA: a power of two up to 32");
    typedef QeGroupCache2<NKEYS, CS, LMAX> Cache;
    extern __shared__ __align__(16) unsigned char qe_gsmem[];
    qe_u64 *s_acc = (qe_u64 *)qe_gsmem;                       // [LMAX][NWORDS][COPIES]
    Cache &gc = *(Cache *)(qe_gsmem + (size_t)LMAX * NWORDS * COPIES * 8);
    const volatile Cache &vc = gc;
    const qe_u32 lane = qe_lane(), tid = threadIdx.x;
    const size_t total = (size_t)NP * p.pt_cap;

    for (qe_u32 q = blockIdx.x; q < (qe_u32)NP; q += gridDim.x) {
        if (__syncthreads_or((tid == 0 && (qe_ld_relaxed_u32(p.g_ctl + 1) & 1u)) ? 1 : 0)) break;   // discarded (pass 1, or another CTA)
        const qe_u32 n = p.pt_cursor[q] < p.pt_cap ? p.pt_cursor[q] : p.pt_cap;
        for (int i = (int)tid; i < CS; i += GT) gc.entry[i] = 0u;
        if (tid == 0) { gc.nlocal = 0u; gc.over = 0u; gc.inserted = 0u; }
        for (int i = (int)tid; i < LMAX * NWORDS * COPIES; i += GT) s_acc[i] = P::word_init((i / COPIES) % NWORDS);
        __syncthreads();
        const qe_u64 *src = p.pt_words + (size_t)q * p.pt_cap;
        qe_u64 rec[R][RW], nxt[PREF ? R : 1][RW];
        // records [i0, i0 + R * GT) of a step: thread t takes 2t, 2t + 1 of every 2 * GT (reads past n stay inside the
        // partition's space -- its size is a multiple of 32 -- and fail the valid bit or the bound)
#define QE_GATHER_LOAD(dst, first)                                                                          \
        _Pragma("unroll") for (int pr = 0; pr < PAIRS; pr++) {                                               \
            const qe_u32 i = (first) + ((qe_u32)pr * GT + tid) * 2u;                                         \
            _Pragma("unroll") for (int w = 0; w < RW; w++) {                                                 \
                if (i < n) qe_gather_load2(src + (size_t)w * total + i, dst[2 * pr][w], dst[2 * pr + 1][w]);  \
                else dst[2 * pr][w] = dst[2 * pr + 1][w] = 0ull;                                             \
            }                                                                                                \
        }
        if (PREF) { QE_GATHER_LOAD(rec, 0u) }
        for (qe_u32 i0 = 0; i0 < n; i0 += (qe_u32)R * GT) {
            if (PREF) { QE_GATHER_LOAD(nxt, i0 + (qe_u32)R * GT) } else { QE_GATHER_LOAD(rec, i0) }
            int lg[R];
            qe_u32 miss = 0;
#pragma unroll
            for (int r = 0; r < R; r++) {
                const bool kept = i0 + ((qe_u32)(r / 2) * GT + tid) * 2u + (qe_u32)(r & 1) < n && (rec[r][0] & QE_PT_VALID) != 0ull;
                qe_u64 k[NKEYS];
#pragma unroll
                for (int i = 0; i < NKEYS; i++) k[i] = rec[r][1 + i];
                const qe_u32 nulls = (qe_u32)(rec[r][0] >> 32) & 0xffu;
                const qe_u64 h = qe_group_hash64<NKEYS>(k, nulls);
                const qe_u32 tag = qe_group_entry_tag(h, nulls);
                const qe_u64 e = *(const volatile qe_u64 *)&gc.entry[((qe_u32)h & (qe_u32)(CS / 2 - 1)) * 2u];
                const qe_u32 e0 = (qe_u32)e, e1 = (qe_u32)(e >> 32);
                const bool m0 = (e0 & QE_GM_MATCH) == tag, m1 = (e1 & QE_GM_MATCH) == tag;
                const qe_u32 pick = m0 ? e0 : e1;
                const int cand = (m0 || m1) ? (int)((pick >> 17) & 0x1fffu) : 0;
                bool hit = m0 || m1;
#pragma unroll
                for (int i = 0; i < NKEYS; i++) hit &= vc.gkeys[cand][i] == k[i];
                lg[r] = (kept && hit) ? cand : -1;
                miss |= (kept && !hit) ? (1u << r) : 0u;
                if ((QE_DEBUG_FLAGS & 512) && rec[r][RW - 1] == 0x1234567ull) s_acc[tid] = rec[r][0];       // loads only
            }
            if (miss && !(QE_DEBUG_FLAGS & 512)) {
#pragma unroll
                for (int r = 0; r < R; r++) {
                    if (!((miss >> r) & 1u)) continue;
                    qe_u64 k[NKEYS];
#pragma unroll
                    for (int i = 0; i < NKEYS; i++) k[i] = rec[r][1 + i];
                    const qe_u32 nulls = (qe_u32)(rec[r][0] >> 32) & 0xffu;
                    const qe_u64 h = qe_group_hash64<NKEYS>(k, nulls);
                    const qe_u32 tag = qe_group_entry_tag(h, nulls);
                    int l = -1;
                    for (qe_u32 d = 2; d < 8u && l < 0; d++) {             // the slots after the home bucket
                        const qe_u32 e = vc.entry[(((qe_u32)h & (qe_u32)(CS / 2 - 1)) * 2u + d) & (qe_u32)(CS - 1)];
                        if ((e & QE_GM_MATCH) != tag) continue;
                        const int cand = (int)((e >> 17) & 0x1fffu);
                        bool same = true;
#pragma unroll
                        for (int i = 0; i < NKEYS; i++) same &= vc.gkeys[cand][i] == k[i];
                        if (same) l = cand;
                    }
                    if (l < 0) l = qe_group_local_slow2<NKEYS, CS, LMAX, false>(p, &gc, k, nulls, h);
                    lg[r] = l;
                }
            }
#pragma unroll
            for (int r = 0; r < R; r++) {
                if (lg[r] >= 0 && !(QE_DEBUG_FLAGS & (256 | 512))) {
                    typename P::Out o;
                    P::record_unpack(rec[r], o);
                    qe_u64 *a = s_acc + (size_t)lg[r] * NWORDS * COPIES + (lane & (qe_u32)(COPIES - 1));
                    P::template acc_atomic<COPIES, true>(a, o, rec[r][0] & 0xffffffffull);     // rows of a partition come in any order
                }
            }
            if (PREF) {
#pragma unroll
                for (int r = 0; r < R; r++)
#pragma unroll
                    for (int w = 0; w < RW; w++) rec[r][w] = nxt[r][w];
            }
        }
#undef QE_GATHER_LOAD
        __syncthreads();
        if (((const volatile qe_u32 &)gc.over) != 0u) {
            if (tid == 0) atomicOr(p.g_ctl + 1, 1u);
            break;
        }
        // the partition's groups are complete: into the table, every slot written by exactly one thread
        if (tid == 0) {
            const qe_u32 ng = gc.nlocal < (qe_u32)LMAX ? gc.nlocal : (qe_u32)LMAX;
            if (atomicAdd(p.g_ctl, ng) + ng > p.g_max_groups) atomicOr(p.g_ctl + 1, 2u);
        }
        for (int i = (int)tid; i < CS; i += GT) {
            const qe_u32 e = gc.entry[i];
            if (!(e & QE_GM_READY)) continue;
            const int l = (int)((e >> 17) & 0x1fffu);
            const qe_u32 nulls = (e >> 9) & 0xffu;
            qe_u64 k[NKEYS];
#pragma unroll
            for (int j = 0; j < NKEYS; j++) k[j] = gc.gkeys[l][j];
            const int gs = qe_group_global_slot<NKEYS, false>(p, k, nulls, qe_group_hash<NKEYS>(k, nulls));
            if (gs < 0) continue;                             // table full: flagged
            for (int w = 0; w < NWORDS; w++) {
                const int kind = P::word_kind(w);
                qe_u64 v = s_acc[((size_t)l * NWORDS + w) * COPIES];
                for (int c = 1; c < COPIES; c++) v = qe_group_combine(kind, v, s_acc[((size_t)l * NWORDS + w) * COPIES + c]);
                p.g_acc[(size_t)gs * NWORDS + w] = v;
            }
        }
        __syncthreads();
    }
}

template <class P>
__device__ __forceinline__ void qe_groupby_atomic_body(const QeParams &p)
{
    constexpr int STEPS = P::STEPS, VEC = P::VEC, NKEYS = P::NKEYS, NWORDS = P::NWORDS;
    constexpr int WARP_ROWS = STEPS * 32 * VEC;
    constexpr int TILE = QE_WARPS * WARP_ROWS;
    const qe_u32 warp = qe_warp(), lane = qe_lane();
    for (qe_u32 tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x) {
        const qe_i64 warp_row0 = (qe_i64)tile * TILE + (qe_i64)warp * WARP_ROWS;
        typename P::Out outs[STEPS][VEC];
        qe_u32 keep[STEPS], inb[STEPS];
        qe_tile_eval<P>(p, warp_row0, outs, keep, inb);
#pragma unroll
        for (int s = 0; s < STEPS; s++) {
#pragma unroll
            for (int j = 0; j < VEC; j++) {
                if (!((keep[s] >> j) & 1u)) continue;
                qe_u64 k[NKEYS];
                qe_u32 nulls;
                P::key_pack(outs[s][j], k, nulls);
                const int gs = qe_group_global_slot<NKEYS>(p, k, nulls, qe_group_hash<NKEYS>(k, nulls));
                if (gs >= 0) {
                    const qe_u64 row = (qe_u64)p.sel_base + (qe_u64)(warp_row0 + (qe_i64)s * 32 * VEC + lane * VEC + j);
                    P::template acc_atomic<1>(p.g_acc + (size_t)gs * NWORDS, outs[s][j], row);
                }
            }
        }
    }
}

// atomics on accumulator words with the accumulators' semantics (Math.min / Math.max: NaN wins, -0.0 < 0.0)
__device__ __forceinline__ void qe_atomic_min_f64(qe_u64 *a, double v)
{
    unsigned long long old = *(volatile unsigned long long *)a;
    for (;;) {
        const double cur = qe_bits_f64((qe_i64)old);
        const double nv = qe_math_min(cur, v);
        if ((unsigned long long)qe_f64_bits(nv) == old) return;
        const unsigned long long seen = atomicCAS((unsigned long long *)a, old, (unsigned long long)qe_f64_bits(nv));
        if (seen == old) return;
        old = seen;
    }
}

__device__ __forceinline__ void qe_atomic_max_f64(qe_u64 *a, double v)
{
    unsigned long long old = *(volatile unsigned long long *)a;
    for (;;) {
        const double cur = qe_bits_f64((qe_i64)old);
        const double nv = qe_math_max(cur, v);
        if ((unsigned long long)qe_f64_bits(nv) == old) return;
        const unsigned long long seen = atomicCAS((unsigned long long *)a, old, (unsigned long long)qe_f64_bits(nv));
        if (seen == old) return;
        old = seen;
    }
}

#endif  // QE_DEVICE_CUH
