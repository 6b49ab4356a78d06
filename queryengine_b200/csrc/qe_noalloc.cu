// qe_noalloc.cu -- qe_agg_column_noalloc (include/qegpu.h): SUM / COUNT / MIN / MAX of one DOUBLE column for callers
// that cannot allocate device memory (code under CUDA graph capture, tight serving loops, harnesses that forbid
// cudaMalloc).  No counterpart in the reference.
//
// Same fixed shape as the global aggregation of the pipelines (qe_agg_body / qe_agg_final_kernel): one partial per
// sub-tile of QE_AGG_SUB_ROWS rows, 64 sub-tiles folded per super-tile by one warp, the super-tiles folded in row
// order -- here by one thread on the device instead of the host, so nothing is copied back either.  The partials, the
// scratch memory of this path, live in ONE file-scope __device__ array of a fixed size: every GPU that loads the library
// holds QE_NOALLOC_SCRATCH_BYTES of it until the process exits.  All calls share it; every word a call reads it wrote
// itself before (nothing has to start at zero), but two calls that overlap on the GPU race for it.
//
// Compiled with -fmad=false like the rest of the library.
#include <cuda_runtime.h>

#include <mutex>
#include <string>

#include "../../include/qegpu.h"
#include "qe_device.cuh"
#include "qe_internal.h"

namespace {

constexpr int SPS = QE_AGG_SUPER_ROWS / QE_AGG_SUB_ROWS;                       // sub-tiles per super-tile
constexpr long long MAX_SUB = (QE_NOALLOC_MAX_ROWS + QE_AGG_SUB_ROWS - 1) / QE_AGG_SUB_ROWS;
constexpr long long MAX_SUPER = (MAX_SUB + SPS - 1) / SPS;
constexpr long long SCRATCH_DOUBLES = (MAX_SUB + MAX_SUPER) * 4;               // [sum, count bits, min, max] each
static_assert(SPS == 64, "one warp folds two sub-tiles per lane");
static_assert(SCRATCH_DOUBLES * 8 == QE_NOALLOC_SCRATCH_BYTES, "the header states the size of the scratch array");

// The scratch memory.  In host code this name is only a stand-in for the symbol: the address kernels get comes from
// cudaGetSymbolAddress (scratch_of below), once per device.
static __device__ __align__(256) double qe_noalloc_scratch[SCRATCH_DOUBLES];

constexpr int ROWS_PER_PASS = QE_BLOCK * 2;                                    // one 16-byte load per thread
constexpr int PASSES = QE_AGG_SUB_ROWS / ROWS_PER_PASS;
static_assert(QE_AGG_SUB_ROWS % (ROWS_PER_PASS * 4) == 0, "four passes in flight");

__device__ __forceinline__ void add_row(QeAcc &a, double v, bool on)
{
    if (on) {
        a.count += 1;
        a.sum += v;
        a.mn = qe_math_min(a.mn, v);
        a.mx = qe_math_max(a.mx, v);
    }
}

// level 1: one CTA per sub-tile (grid-stride); thread t takes rows 2t, 2t+1 of every pass of 512 rows, passes in
// ascending order, then the lane tree and the warps in sequence (qe_acc_block_reduce_store_all)
__global__ void __launch_bounds__(QE_BLOCK) qe_noalloc_sub_kernel(const double *__restrict__ values, const qe_u8 *__restrict__ valid,
                                                                   qe_i64 nrows, qe_i64 nsub, double *sub)
{
    __shared__ QeAcc s_slot[QE_WARPS];
    for (qe_i64 s = blockIdx.x; s < nsub; s += gridDim.x) {
        QeAcc a;
        qe_acc_init(a);
        const qe_i64 base = s * QE_AGG_SUB_ROWS + 2 * (qe_i64)threadIdx.x;
        for (int p0 = 0; p0 < PASSES; p0 += 4) {
            double2 v[4];
            qe_u32 on[4];
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const qe_i64 r = base + (qe_i64)(p0 + j) * ROWS_PER_PASS;
                v[j] = make_double2(0.0, 0.0);
                on[j] = 0u;
                if (r + 1 < nrows) {
                    v[j] = *reinterpret_cast<const double2 *>(values + r);
                    on[j] = 3u;
                } else if (r < nrows) {
                    v[j].x = values[r];
                    on[j] = 1u;
                }
                if (valid && on[j]) on[j] &= (qe_u32)(valid[r >> 3] >> (r & 7));      // r is even: both bits in one byte
            }
#pragma unroll
            for (int j = 0; j < 4; j++) {
                add_row(a, v[j].x, on[j] & 1u);
                add_row(a, v[j].y, on[j] & 2u);
            }
        }
        qe_acc_block_reduce_store_all(a, s_slot, sub, s, 1, 0);
        __syncthreads();                                                          // s_slot is reused by the next sub-tile
    }
}

// level 2: one warp per super-tile, lane l takes sub-tiles 2l and 2l+1, then the lane tree (the shape of
// qe_agg_final_kernel in qe_kernels.cu)
__global__ void __launch_bounds__(QE_BLOCK) qe_noalloc_super_kernel(const double *sub, qe_i64 nsub, double *sup)
{
    const qe_i64 super = (qe_i64)blockIdx.x * QE_WARPS + qe_warp();
    if (super * SPS >= nsub) return;
    const qe_u32 lane = qe_lane();
    QeAcc a;
    qe_acc_init(a);
#pragma unroll
    for (int h = 0; h < 2; h++) {
        const qe_i64 i = super * SPS + 2 * (qe_i64)lane + h;
        if (i < nsub) {
            QeAcc x;
            x.sum = sub[i * 4 + 0];
            x.count = qe_f64_bits(sub[i * 4 + 1]);
            x.mn = sub[i * 4 + 2];
            x.mx = sub[i * 4 + 3];
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
        sup[super * 4 + 0] = a.sum;
        sup[super * 4 + 1] = qe_bits_f64(a.count);
        sup[super * 4 + 2] = a.mn;
        sup[super * 4 + 3] = a.mx;
    }
}

// level 3: the super-tiles in row order (what fold_aggs in qe_runtime.cu does on the host), into the caller's buffer
__global__ void qe_noalloc_fold_kernel(const double *sup, qe_i64 nsuper, double *out)
{
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    QeAcc a;
    qe_acc_init(a);
    for (qe_i64 i = 0; i < nsuper; i++) {
        QeAcc x;
        x.sum = sup[i * 4 + 0];
        x.count = qe_f64_bits(sup[i * 4 + 1]);
        x.mn = sup[i * 4 + 2];
        x.mx = sup[i * 4 + 3];
        qe_acc_merge_all(a, x);
    }
    out[0] = a.sum;
    out[1] = (double)a.count;
    out[2] = a.mn;
    out[3] = a.mx;
}

constexpr int MAX_DEVICES = 64;
std::mutex g_mu;
double *g_scratch[MAX_DEVICES];          // device address of qe_noalloc_scratch on every device seen so far
int g_sms[MAX_DEVICES];

// set-up, once per device, with that device current
int scratch_of(int device, double **ptr, int *sms)
{
    std::lock_guard<std::mutex> g(g_mu);
    if (!g_scratch[device]) {
        void *p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, qe_noalloc_scratch);
        if (e != cudaSuccess || !p)
            return qe_fail(QE_ERR_CUDA, std::string("cudaGetSymbolAddress(qe_noalloc_scratch): ") + cudaGetErrorString(e));
        int n = 0;
        e = cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, device);
        if (e != cudaSuccess || n <= 0) return qe_fail(QE_ERR_CUDA, std::string("cudaDeviceGetAttribute: ") + cudaGetErrorString(e));
        g_sms[device] = n;
        g_scratch[device] = (double *)p;
    }
    *ptr = g_scratch[device];
    *sms = g_sms[device];
    return QE_OK;
}

struct DeviceGuard {
    int prev = -1;
    bool switched = false;
    ~DeviceGuard() { if (switched) cudaSetDevice(prev); }
};

}  // namespace

extern "C" int qe_agg_column_noalloc(int device, const double *d_values, const uint8_t *d_validity, int64_t nrows,
                                     double *d_out, void *stream)
{
    if (device < 0 || device >= MAX_DEVICES) return qe_fail(QE_ERR_INVALID_ARGUMENT, "qe_agg_column_noalloc: bad device index");
    if (nrows < 0) return qe_fail(QE_ERR_INVALID_ARGUMENT, "qe_agg_column_noalloc: nrows is negative");
    if (nrows > (int64_t)QE_NOALLOC_MAX_ROWS)
        return qe_fail(QE_ERR_INVALID_ARGUMENT, "qe_agg_column_noalloc: " + std::to_string(nrows) + " rows exceed QE_NOALLOC_MAX_ROWS = " +
                                                    std::to_string((long long)QE_NOALLOC_MAX_ROWS) +
                                                    " (the scratch array has a fixed size; split the column or use qe_pipeline_run)");
    if (!d_out || (nrows > 0 && !d_values)) return qe_fail(QE_ERR_INVALID_ARGUMENT, "qe_agg_column_noalloc: d_values / d_out is NULL");
    if (((uintptr_t)d_values & 15u) || ((uintptr_t)d_out & 7u))
        return qe_fail(QE_ERR_INVALID_ARGUMENT, "qe_agg_column_noalloc: d_values must be 16-byte aligned, d_out 8-byte aligned");
    DeviceGuard guard;
    cudaError_t e = cudaGetDevice(&guard.prev);
    if (e != cudaSuccess) return qe_fail(QE_ERR_CUDA, std::string("cudaGetDevice: ") + cudaGetErrorString(e));
    if (guard.prev != device) {
        e = cudaSetDevice(device);
        if (e != cudaSuccess) return qe_fail(QE_ERR_CUDA, std::string("cudaSetDevice: ") + cudaGetErrorString(e));
        guard.switched = true;
    }
    double *scratch = nullptr;
    int sms = 0;
    const int rc = scratch_of(device, &scratch, &sms);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    const int64_t nsub = (nrows + QE_AGG_SUB_ROWS - 1) / QE_AGG_SUB_ROWS;
    const int64_t nsuper = (nsub + SPS - 1) / SPS;
    double *sub = scratch, *sup = scratch + MAX_SUB * 4;
    if (nsub > 0) {
        const int64_t grid = nsub < (int64_t)sms * 8 ? nsub : (int64_t)sms * 8;
        qe_noalloc_sub_kernel<<<(unsigned)grid, QE_BLOCK, 0, st>>>(d_values, d_validity, (qe_i64)nrows, (qe_i64)nsub, sub);
        qe_noalloc_super_kernel<<<(unsigned)((nsuper + QE_WARPS - 1) / QE_WARPS), QE_BLOCK, 0, st>>>(sub, (qe_i64)nsub, sup);
    }
    qe_noalloc_fold_kernel<<<1, 32, 0, st>>>(sup, (qe_i64)nsuper, d_out);
    e = cudaGetLastError();
    if (e != cudaSuccess) return qe_fail(QE_ERR_CUDA, std::string("qe_agg_column_noalloc launch: ") + cudaGetErrorString(e));
    return QE_OK;
}
