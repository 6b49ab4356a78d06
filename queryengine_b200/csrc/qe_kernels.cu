// qe_kernels.cu -- ahead-of-time compiled kernels of libqegpu.so (sm_100a):
//   * the three kernel bodies of qe_device.cuh instantiated with the generic
//     interpreter program (always-available path, cross-check for generated code)
//   * the final fixed-order reduce of per-CTA aggregate partials (K4's on-GPU half)
//   * the synthetic column generators (K5)
// Compiled with -fmad=false so arithmetic matches the JVM's unfused DMUL/DADD.
#include <cuda_runtime.h>

#include "qe_interp.cuh"
#include "qe_kernels.h"

namespace {

using InterpProgram = QeInterpProgram<QE_INTERP_STEPS, QE_INTERP_VEC>;

__global__ void __launch_bounds__(QE_BLOCK) qe_interp_filter_kernel(const __grid_constant__ QeParams p)
{
    qe_filter_body<InterpProgram>(p);
}

__global__ void __launch_bounds__(QE_BLOCK) qe_interp_project_kernel(const __grid_constant__ QeParams p)
{
    qe_project_body<InterpProgram>(p);
}

__global__ void __launch_bounds__(QE_BLOCK) qe_interp_agg_kernel(const __grid_constant__ QeParams p)
{
    qe_agg_body<InterpProgram>(p);
}

// Second level of the fixed-shape aggregation (QE_AGG_SUPER_ROWS in qe_device.cuh): one warp per super-tile folds
// its 64 consecutive sub-tile partials -- lane l takes sub-tiles 2l and 2l+1, then the lane tree.  A missing
// sub-tile (past the end of the batch) is the identity; x + 0.0 is exact and a partial sum is never -0.0 (sums
// start at +0.0, M/evaluator/Accumulators.kt:40), so padding does not change a bit.
__global__ void __launch_bounds__(QE_BLOCK) qe_agg_final_kernel(const double *sub, qe_i64 nsub, int nagg, double *out)
{
    constexpr int SPS = QE_AGG_SUPER_ROWS / QE_AGG_SUB_ROWS;
    static_assert(SPS == 64, "one warp folds two sub-tiles per lane");
    const qe_i64 super = (qe_i64)blockIdx.x * QE_WARPS + qe_warp();
    if (super * SPS >= nsub) return;
    const qe_u32 lane = qe_lane();
    for (int k = 0; k < nagg; k++) {
        QeAcc a;
        qe_acc_init(a);
#pragma unroll
        for (int h = 0; h < 2; h++) {
            const qe_i64 i = super * SPS + 2 * (qe_i64)lane + h;
            if (i < nsub) {
                const double *src = sub + (i * nagg + k) * 4;
                QeAcc x;
                x.sum = src[0];
                x.count = qe_f64_bits(src[1]);
                x.mn = src[2];
                x.mx = src[3];
                qe_acc_merge_all(a, x);
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            QeAcc b;
            b.count = qe_shfl_down_i64(a.count, o);
            b.sum = qe_shfl_down_f64(a.sum, o);
            b.mn = qe_shfl_down_f64(a.mn, o);
            b.mx = qe_shfl_down_f64(a.mx, o);
            qe_acc_merge_all(a, b);
        }
        if (lane == 0) {
            double *dst = out + (super * nagg + k) * 4;
            dst[0] = a.sum;
            dst[1] = qe_bits_f64(a.count);
            dst[2] = a.mn;
            dst[3] = a.mx;
        }
    }
}

// ---- synthetic columns ------------------------------------------------------
// Counter-based: value(i) is a pure function of (seed, global row index i).

__device__ __forceinline__ qe_u64 qe_splitmix64(qe_u64 x)
{
    x += 0x9e3779b97f4a7c15ull;
    x = (x ^ (x >> 30)) * 0xbf58476d1ce4e5b9ull;
    x = (x ^ (x >> 27)) * 0x94d049bb133111ebull;
    return x ^ (x >> 31);
}

__device__ __forceinline__ double qe_gen_value(int kind, qe_u64 seed, qe_u64 i, double p0, double p1)
{
    const qe_u64 h = qe_splitmix64(seed ^ i);
    switch (kind) {
    case 0: return (double)(h >> 24);                                          // QE_GEN_UNIFORM_U40
    case 1: {                                                                  // QE_GEN_UNIFORM_INT
        const qe_u64 span = (qe_u64)(p1 - p0) + 1ull;
        return p0 + (double)(h % span);
    }
    case 2: return p0 + ((double)(h >> 11) * (1.0 / 9007199254740992.0)) * (p1 - p0);   // QE_GEN_UNIFORM_REAL
    case 3: {                                                                  // QE_GEN_CENTS
        const qe_u64 span = (qe_u64)(p1 - p0) + 1ull;
        return (p0 + (double)(h % span)) / 100.0;
    }
    default: return (double)i;                                                 // QE_GEN_ROW_INDEX
    }
}

__global__ void qe_generate_kernel(void *values, qe_u32 *validity, int type, int kind, qe_u64 seed, double p0, double p1,
                                   double null_rate, qe_i64 row_offset, qe_i64 nrows)
{
    // one thread per 32 rows so that validity / boolean words are written whole
    const qe_i64 word = (qe_i64)blockIdx.x * blockDim.x + threadIdx.x;
    const qe_i64 r0 = word * 32;
    if (r0 >= nrows) return;
    qe_u32 vbits = 0, bbits = 0;
    const qe_u64 null_threshold = null_rate >= 1.0 ? ~0ull : (qe_u64)(null_rate * 18446744073709551616.0);
    for (int j = 0; j < 32; j++) {
        const qe_i64 r = r0 + j;
        if (r >= nrows) break;
        const qe_u64 g = (qe_u64)(row_offset + r);
        const double v = qe_gen_value(kind, seed, g, p0, p1);
        if (type == QE_T_F64) ((double *)values)[r] = v;
        else if (type == QE_T_I64) ((qe_i64 *)values)[r] = (qe_i64)v;
        else if (type == QE_T_DATE32 || type == QE_T_STR) ((qe_i32 *)values)[r] = (qe_i32)v;
        else bbits |= (((qe_u64)v) & 1ull ? 1u : 0u) << j;
        const bool valid = qe_splitmix64(~seed ^ (g * 0x9e3779b97f4a7c15ull)) >= null_threshold || null_rate <= 0.0;
        vbits |= (valid ? 1u : 0u) << j;
    }
    if (type == QE_T_BOOL) ((qe_u32 *)values)[word] = bbits;
    if (validity) validity[word] = vbits;
}


// ---- yardsticks ---------------------------------------------------------------
// What this device's HBM does for a pure read, a 1:1 copy and a pure write, measured with the access shapes the
// pipelines use (256-bit loads / stores, evict-first): the denominators of bench.py's roofline for a launch's own
// read / write mix (a filter at 1 % selectivity is 99 % reads, at 99 % it writes more than it reads).
__global__ void __launch_bounds__(256) qe_yardstick_kernel(const qe_i64 *src, qe_i64 *dst, qe_i64 n4, int kind, qe_i64 *sink)
{
    qe_i64 acc = 0;
    const qe_i64 stride = (qe_i64)gridDim.x * blockDim.x;
    for (qe_i64 i = (qe_i64)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
        qe_i64 a = i, b = i, c = i, d = i;
        if (kind != 2) qe_ldg_i64x4(src + 4 * i, a, b, c, d);
        if (kind == 0) acc ^= a ^ b ^ c ^ d;
        else asm volatile("st.global.v4.s64 [%0], {%1,%2,%3,%4};" :: "l"(dst + 4 * i), "l"(a), "l"(b), "l"(c), "l"(d) : "memory");
    }
    if (kind == 0 && acc == 0x5eed5eed5eedLL) *sink = acc;      // keeps the loads alive
}

// ---- group-by helpers (the main kernels are generated; see "group-by" in qe_device.cuh)

// every slot empty, every accumulator word at its initial value
__global__ void qe_groupby_init_kernel(qe_u32 *state, qe_u64 *acc, qe_u32 *ctl, qe_i64 nslots, qe::GroupWords gw)
{
    const qe_i64 i = (qe_i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0) { ctl[0] = 0u; ctl[1] = 0u; ctl[2] = 0u; ctl[3] = 0u; }
    if (i >= nslots) return;
    state[i] = 0u;
    for (int w = 0; w < gw.nwords; w++) acc[i * gw.nwords + w] = gw.init[w];
}

// Private variant: fold the per-CTA partials into the table, CTA 0 first, then CTA 1, ... (one CTA of threads;
// within one source CTA the local groups map to distinct slots, so there are no conflicts and the order of
// the additions into a slot is the CTA order: reproducible).
// Few groups: one CTA.  Many groups (hundreds to thousands per source CTA): several CTAs, CTA b takes the table slots
// with slot % gridDim.x == b -- every one of them still sees its slots' partials in source-CTA order, so the result
// has the same bits for any number of merging CTAs (one CTA alone took ~1 us per group: 3 ms for 3 000 groups).
__global__ void qe_groupby_merge_kernel(const qe_u64 *part, const qe_u32 *part_slot, qe_u64 *acc, int nctas, int lmax,
                                        qe::GroupWords gw)
{
    const qe_u32 nb = gridDim.x, b = blockIdx.x;
    for (int c = 0; c < nctas; c++) {
        const qe_u32 *ps = part_slot + (size_t)c * (1 + lmax);
        const int n = (int)ps[0];
        if (nb == 1u) {
            for (int idx = (int)threadIdx.x; idx < n * gw.nwords; idx += (int)blockDim.x) {
                const int l = idx / gw.nwords, w = idx - l * gw.nwords;
                const qe_u32 slot = ps[1 + l];
                if (slot == 0xffffffffu) continue;
                qe_u64 *dst = acc + (size_t)slot * gw.nwords + w;
                *dst = qe_group_combine(gw.kind[w], *dst, part[((size_t)c * lmax + l) * gw.nwords + w]);
            }
        } else {
            for (int l = (int)threadIdx.x; l < n; l += (int)blockDim.x) {
                const qe_u32 slot = ps[1 + l];
                if (slot == 0xffffffffu || slot % nb != b) continue;
                for (int w = 0; w < gw.nwords; w++) {
                    qe_u64 *dst = acc + (size_t)slot * gw.nwords + w;
                    *dst = qe_group_combine(gw.kind[w], *dst, part[((size_t)c * lmax + l) * gw.nwords + w]);
                }
            }
        }
        __syncthreads();
    }
}

// compact the occupied slots into records [acc words][key words][nulls], in no particular order
__global__ void qe_groupby_collect_kernel(const qe_u32 *state, const qe_u64 *keys, const qe_u32 *nulls, const qe_u64 *acc,
                                          qe_i64 nslots, int nkeys, int nwords, qe_u64 *records, qe_u32 *ctl, qe_u32 capacity)
{
    const qe_i64 i = (qe_i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nslots || state[i] != 2u) return;
    const qe_u32 at = atomicAdd(ctl + 2, 1u);
    if (at >= capacity) return;
    qe_u64 *r = records + (size_t)at * (nwords + nkeys + 1);
    for (int w = 0; w < nwords; w++) r[w] = acc[i * nwords + w];
    for (int k = 0; k < nkeys; k++) r[nwords + k] = keys[i * nkeys + k];
    r[nwords + nkeys] = nulls[i];
}

}  // namespace

namespace qe {

cudaError_t launch_interp(int mode, const QeParams &p, int grid, cudaStream_t stream)
{
    switch (mode) {
    case 0: qe_interp_filter_kernel<<<grid, QE_BLOCK, 0, stream>>>(p); break;
    case 1: qe_interp_project_kernel<<<grid, QE_BLOCK, 0, stream>>>(p); break;
    default: qe_interp_agg_kernel<<<grid, QE_BLOCK, 0, stream>>>(p); break;
    }
    return cudaGetLastError();
}

int interp_blocks_per_sm(int mode)
{
    int n = 0;
    const void *f = mode == 0 ? (const void *)qe_interp_filter_kernel
                  : mode == 1 ? (const void *)qe_interp_project_kernel : (const void *)qe_interp_agg_kernel;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, f, QE_BLOCK, 0) != cudaSuccess || n < 1) n = 1;
    return n;
}

__global__ void qe_store_u64_kernel(unsigned long long *dst, const unsigned long long *src)
{
    *dst = *src;
    __threadfence_system();
}

cudaError_t launch_store_u64(unsigned long long *dst, const unsigned long long *src, cudaStream_t stream)
{
    qe_store_u64_kernel<<<1, 1, 0, stream>>>(dst, src);
    return cudaGetLastError();
}

cudaError_t launch_agg_final(const double *sub_partials, int64_t nsub, int nagg, double *super_partials, cudaStream_t stream)
{
    const int64_t sps = QE_AGG_SUPER_ROWS / QE_AGG_SUB_ROWS;
    const int64_t nsuper = (nsub + sps - 1) / sps;
    if (nsuper <= 0) return cudaSuccess;
    qe_agg_final_kernel<<<(unsigned)((nsuper + QE_WARPS - 1) / QE_WARPS), QE_BLOCK, 0, stream>>>(sub_partials, (qe_i64)nsub, nagg, super_partials);
    return cudaGetLastError();
}

cudaError_t launch_groupby_init(uint32_t *state, uint64_t *acc, uint32_t *ctl, int64_t nslots, const GroupWords &gw, cudaStream_t stream)
{
    const int block = 256;
    qe_groupby_init_kernel<<<(unsigned)((nslots + block - 1) / block), block, 0, stream>>>(state, (qe_u64 *)acc, ctl, (qe_i64)nslots, gw);
    return cudaGetLastError();
}

cudaError_t launch_groupby_merge(const uint64_t *part, const uint32_t *part_slot, uint64_t *acc, int nctas, int lmax,
                                 const GroupWords &gw, cudaStream_t stream)
{
    qe_groupby_merge_kernel<<<lmax >= 256 ? 64 : 1, 256, 0, stream>>>((const qe_u64 *)part, part_slot, (qe_u64 *)acc, nctas, lmax, gw);
    return cudaGetLastError();
}

cudaError_t launch_groupby_collect(const uint32_t *state, const uint64_t *keys, const uint32_t *nulls, const uint64_t *acc,
                                   int64_t nslots, int nkeys, int nwords, uint64_t *records, uint32_t *ctl, uint32_t capacity,
                                   cudaStream_t stream)
{
    const int block = 256;
    qe_groupby_collect_kernel<<<(unsigned)((nslots + block - 1) / block), block, 0, stream>>>(
        state, (const qe_u64 *)keys, nulls, (const qe_u64 *)acc, (qe_i64)nslots, nkeys, nwords, (qe_u64 *)records, ctl, capacity);
    return cudaGetLastError();
}

cudaError_t launch_yardstick(const void *src, void *dst, int64_t bytes, int kind, int sms, void *sink, cudaStream_t stream)
{
    qe_yardstick_kernel<<<sms * 8, 256, 0, stream>>>((const qe_i64 *)src, (qe_i64 *)dst, (qe_i64)(bytes / 32), kind, (qe_i64 *)sink);
    return cudaGetLastError();
}

cudaError_t launch_generate(void *values, uint32_t *validity, int type, int kind, uint64_t seed, double p0, double p1,
                            double null_rate, int64_t row_offset, int64_t nrows, cudaStream_t stream)
{
    const int64_t words = (nrows + 31) / 32;
    if (words == 0) return cudaSuccess;
    const int block = 128;
    const int64_t grid = (words + block - 1) / block;
    qe_generate_kernel<<<(unsigned)grid, block, 0, stream>>>(values, (qe_u32 *)validity, type, kind, (qe_u64)seed, p0, p1,
                                                             null_rate, (qe_i64)row_offset, (qe_i64)nrows);
    return cudaGetLastError();
}

}  // namespace qe
