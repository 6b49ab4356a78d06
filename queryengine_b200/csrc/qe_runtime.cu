// qe_runtime.cu -- host runtime behind the C ABI of include/qegpu.h: context,
// device memory pool, columnar tables with pinned-staged asynchronous uploads,
// pipeline binding / specialisation / launch, results.  Host code only; the
// kernels are in qe_device.cuh (bodies), qe_kernels.cu (ahead of time) and the
// NVRTC-compiled output of qe_codegen.cpp.
//
// There is no CPU evaluation path anywhere in this file: a pipeline either runs
// on the GPU or the call fails.
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <list>
#include <map>
#include <mutex>
#include <string>
#include <unordered_map>
#include <vector>

#include "../../include/qegpu.h"
#include "qe_codegen.h"
#include "qe_cubin_cache.h"
#include "qe_device.cuh"
#include "qe_internal.h"
#include "qe_interp.cuh"
#include "qe_jit.h"
#include "qe_kernels.h"
#include "qe_program.h"

#define QE_VERSION_STRING "0.2.0"

// Column allocations are padded so that any tile shape (<= 8192 rows) starting
// at any 128-aligned batch offset can be loaded without bounds checks.
static const int64_t kPadRows = 8192;
static const size_t kStagingBytes = 16u << 20;

// ------------------------------------------------------------------ errors

static thread_local std::string tls_error;

static int fail(int code, const std::string &msg)
{
    tls_error = msg;
    return code;
}

int qe_fail(int code, const std::string &msg) { return fail(code, msg); }

static int fail_cuda(const char *what, cudaError_t e)
{
    tls_error = std::string(what) + ": " + cudaGetErrorName(e) + " (" + cudaGetErrorString(e) + ")";
    return QE_ERR_CUDA;
}

#define QE_CUDA(call) do { cudaError_t e__ = (call); if (e__ != cudaSuccess) return fail_cuda(#call, e__); } while (0)

extern "C" const char *qe_last_error(void) { return tls_error.c_str(); }
extern "C" const char *qe_version(void) { return QE_VERSION_STRING; }

// ------------------------------------------------------------- memory pool
// Freed blocks are kept by size and handed out again, so steady-state pipeline
// runs do no cudaMalloc / cudaFree (both synchronise the device).

struct Pool {
    std::multimap<size_t, void *> free_blocks;
    std::unordered_map<void *, size_t> sizes;
    size_t live = 0, reserved = 0;

    static size_t round(size_t n)
    {
        const size_t q = n < (1u << 20) ? 512 : (2u << 20);
        return (n + q - 1) / q * q;
    }

    cudaError_t alloc(size_t n, void **out)
    {
        n = round(n ? n : 1);
        auto it = free_blocks.lower_bound(n);
        if (it != free_blocks.end() && it->first <= n + n / 4) {
            *out = it->second;
            free_blocks.erase(it);
            live += sizes[*out];
            return cudaSuccess;
        }
        cudaError_t e = cudaMalloc(out, n);
        if (e != cudaSuccess) {
            // give cached blocks back and retry once
            release_cached();
            e = cudaMalloc(out, n);
            if (e != cudaSuccess) return e;
        }
        sizes[*out] = n;
        reserved += n;
        live += n;
        return cudaSuccess;
    }

    void free(void *p)
    {
        if (!p) return;
        auto it = sizes.find(p);
        if (it == sizes.end()) return;
        live -= it->second;
        free_blocks.emplace(it->second, p);
    }

    void release_cached()
    {
        for (auto &kv : free_blocks) {
            cudaFree(kv.second);
            reserved -= kv.first;
            sizes.erase(kv.second);
        }
        free_blocks.clear();
    }

    void destroy()
    {
        release_cached();
        for (auto &kv : sizes) cudaFree(kv.first);
        sizes.clear();
    }
};

// ----------------------------------------------------------------- handles

struct JitEntry {
    qe::JitKernel kernel;
    qe::TileShape shape;
    std::list<std::string>::iterator lru;      // position in qe_ctx::jit_lru
};

struct qe_ctx {
    int device = 0;
    int num_sms = 0;
    cudaStream_t stream = nullptr;
    cudaStream_t h2d = nullptr, d2h = nullptr;   // copy streams of qe_pipeline_stream (created on first use)
    Pool pool;
    // kernels loaded on this device, by pipeline signature + tile shape; least recently used first in jit_lru.
    // (The compiled cubins themselves live in the process-wide store of qe_cubin_cache.cpp.)
    std::unordered_map<std::string, JitEntry> jit_cache;
    std::list<std::string> jit_lru;
    // what a filter kept the last time it ran in this context, by query (signature + literal values): the shape hint of a
    // run that has no earlier result of its own to look at (a new operator running a query the context has seen)
    std::unordered_map<std::string, double> sel_memo;
    int64_t opt_jit_cache_entries = 128;
    int64_t jit_loads = 0;                       // cudaLibraryLoadData calls (cold or evicted kernels)
    // options
    int64_t opt_jit = 1, opt_steps = 0, opt_vec = 0, opt_grid_per_sm = 0, opt_staging = -1, opt_stages = 0, opt_debug_flags = 0, opt_pipeline = 1, opt_groupby_registers = 1, opt_groupby_lanes = 1, opt_lane_acc_kb = 0;
    int64_t opt_min_blocks = 0;
    int64_t opt_groupby_twopass = 1;    // more groups than a CTA holds: partition the rows by key hash before global atomics
    int64_t opt_groupby_warps = 0;      // one copy per warp: measured slower than the CAS loops, on request only
    int64_t opt_nlb = 0, opt_loads = 0, opt_tile_rows = 0, opt_ncw = 0, opt_defer = 0, opt_stage_kb = 0;
    int64_t opt_list_cap = 0;                                // ring filter, sparse tiles: 0 auto (32), -1 off, 32 / 64 / 96 / 128
    int64_t opt_nsw = 0;                                     // ... sweeper warps: 0 auto (2), -1 none (consumers store their own lists), 1..4
    int64_t launches = 0;
    bool warned_no_nvrtc = false;
    // pinned staging ring for uploads from pageable memory
    void *staging[2] = {nullptr, nullptr};
    cudaEvent_t staging_free[2] = {nullptr, nullptr};
};

struct qe_program {
    qe_ctx *ctx;
    qe::Program prog;
};

struct qe_table {
    qe_ctx *ctx;
    int64_t nrows = 0;
    int64_t padded_rows = 0;
    std::vector<qe_col_desc> cols;
    std::vector<void *> values;
    std::vector<uint32_t *> valid;
};

static size_t type_width(int type)
{
    switch (type) {
    case QE_TYPE_DOUBLE: case QE_TYPE_INT64: return 8;
    case QE_TYPE_DATE32: case QE_TYPE_STRING: return 4;      // STRING = int32 dictionary codes
    default: return 0;    // BOOLEAN: bitmap
    }
}

static size_t column_bytes(int type, int64_t rows)
{
    if (type == QE_TYPE_BOOLEAN) return (size_t)((rows + 31) / 32) * 4;
    return (size_t)rows * type_width(type);
}

struct DeviceGuard {
    int prev = -1;
    explicit DeviceGuard(int dev)
    {
        cudaGetDevice(&prev);
        if (prev != dev) cudaSetDevice(dev); else prev = -1;
    }
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

// ----------------------------------------------------------------- context

extern "C" int qe_ctx_create(int device, qe_ctx **out)
{
    if (!out) return fail(QE_ERR_INVALID_ARGUMENT, "out is NULL");
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) {
        (void)cudaGetLastError();
        return fail(QE_ERR_CUDA, std::string("no CUDA device available (") + cudaGetErrorString(e) +
                                 "); libqegpu has no CPU fallback");
    }
    if (device < 0 || device >= count) return fail(QE_ERR_INVALID_ARGUMENT, "device index out of range");
    QE_CUDA(cudaSetDevice(device));
    cudaDeviceProp prop;
    QE_CUDA(cudaGetDeviceProperties(&prop, device));
    if (prop.major < 10) {
        return fail(QE_ERR_UNSUPPORTED, std::string("device '") + prop.name + "' is sm_" + std::to_string(prop.major) +
                                        std::to_string(prop.minor) + "; this library is built for sm_100a (B200) only");
    }
    qe_ctx *ctx = new qe_ctx();
    ctx->device = device;
    ctx->num_sms = prop.multiProcessorCount;
    e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking);
    if (e != cudaSuccess) { delete ctx; return fail_cuda("cudaStreamCreate", e); }
    *out = ctx;
    return QE_OK;
}

extern "C" int qe_ctx_destroy(qe_ctx *ctx)
{
    if (!ctx) return QE_OK;
    DeviceGuard g(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    for (auto &kv : ctx->jit_cache) qe::jit_unload(kv.second.kernel);
    for (int i = 0; i < 2; i++) {
        if (ctx->staging[i]) cudaFreeHost(ctx->staging[i]);
        if (ctx->staging_free[i]) cudaEventDestroy(ctx->staging_free[i]);
    }
    ctx->pool.destroy();
    if (ctx->h2d) cudaStreamDestroy(ctx->h2d);
    if (ctx->d2h) cudaStreamDestroy(ctx->d2h);
    cudaStreamDestroy(ctx->stream);
    delete ctx;
    return QE_OK;
}

extern "C" int qe_ctx_synchronize(qe_ctx *ctx)
{
    if (!ctx) return fail(QE_ERR_INVALID_ARGUMENT, "ctx is NULL");
    DeviceGuard g(ctx->device);
    QE_CUDA(cudaStreamSynchronize(ctx->stream));
    return QE_OK;
}

extern "C" void *qe_ctx_stream(qe_ctx *ctx) { return ctx ? (void *)ctx->stream : nullptr; }

extern "C" int qe_ctx_set_option(qe_ctx *ctx, const char *key, int64_t value)
{
    if (!ctx || !key) return fail(QE_ERR_INVALID_ARGUMENT, "ctx or key is NULL");
    const std::string k(key);
    if (k == "jit") ctx->opt_jit = value;
    else if (k == "steps") {
        if (value != 0 && value != 1 && value != 2 && value != 4 && value != 8)
            return fail(QE_ERR_INVALID_ARGUMENT, "steps must be 0 (auto), 1, 2, 4 or 8");
        ctx->opt_steps = value;
    } else if (k == "vec") {
        if (value != 0 && value != 2 && value != 4) return fail(QE_ERR_INVALID_ARGUMENT, "vec must be 0 (auto), 2 or 4");
        ctx->opt_vec = value;
    } else if (k == "grid_per_sm") {
        if (value < 0 || value > 32) return fail(QE_ERR_INVALID_ARGUMENT, "grid_per_sm must be in [0, 32]");
        ctx->opt_grid_per_sm = value;
    } else if (k == "staging") {
        if (value < -1 || value > 2) return fail(QE_ERR_INVALID_ARGUMENT, "staging must be -1 (auto), 0, 1 or 2");
        ctx->opt_staging = value;
    } else if (k == "min_blocks") {
        if (value < 0 || value > 8) return fail(QE_ERR_INVALID_ARGUMENT, "min_blocks must be in [0, 8]");
        ctx->opt_min_blocks = value;
    } else if (k == "pipeline") {
        ctx->opt_pipeline = value != 0;
    } else if (k == "lane_acc_kb") {
        if (value < 0 || value > 200) return fail(QE_ERR_INVALID_ARGUMENT, "lane_acc_kb must be in [0, 200] (0 = no cap)");
        ctx->opt_lane_acc_kb = value;
    } else if (k == "groupby_lanes") {
        ctx->opt_groupby_lanes = value != 0;                 // 0: skip the lane-copies variant (32 copies per CTA)
    } else if (k == "groupby_twopass") {
        ctx->opt_groupby_twopass = value != 0;               // 0: more groups than a CTA's shared memory holds go straight to global atomics
    } else if (k == "groupby_warps") {
        ctx->opt_groupby_warps = value != 0;                 // 1 tries the one-copy-per-warp variant before the single-copy one
    } else if (k == "groupby_registers") {
        if (value < 0 || value > 2) return fail(QE_ERR_INVALID_ARGUMENT, "groupby_registers must be 0 (off), 1 (by the cost rule) or 2 (whenever the registers suffice)");
        ctx->opt_groupby_registers = value;                  // 0: group-by starts with the private (shared-memory) variant
    } else if (k == "tile_rows") {
        if (value < 0 || value > kPadRows || value % 128) return fail(QE_ERR_INVALID_ARGUMENT, "tile_rows must be 0 (auto) or a multiple of 128 up to 8192");
        ctx->opt_tile_rows = value;
    } else if (k == "ncw") {
        if (value < 0 || value > 28 || (value % 4 && value != 11))
            return fail(QE_ERR_INVALID_ARGUMENT, "ncw must be 0 (auto), a multiple of 4 up to 28, or 11 (group-by register variant)");
        ctx->opt_ncw = value;
    } else if (k == "defer") {
        if (value < 0 || value > 30) return fail(QE_ERR_INVALID_ARGUMENT, "defer must be in [0, 30] (0 = auto)");
        ctx->opt_defer = value;
    } else if (k == "nlb") {
        if (value < 0 || value > 8) return fail(QE_ERR_INVALID_ARGUMENT, "nlb must be in [0, 8] (0 = auto)");
        ctx->opt_nlb = value;
    } else if (k == "loads") {
        if (value < 0 || value > 32) return fail(QE_ERR_INVALID_ARGUMENT, "loads must be in [0, 32] (0 = no limit)");
        ctx->opt_loads = value;
    } else if (k == "list_cap") {
        if (value != -1 && (value < 0 || value > 128 || value % 32))
            return fail(QE_ERR_INVALID_ARGUMENT, "list_cap must be 0 (auto), -1 (off) or 32, 64, 96, 128");
        ctx->opt_list_cap = value;
    } else if (k == "nsw") {
        if (value < -1 || value > 4) return fail(QE_ERR_INVALID_ARGUMENT, "nsw must be 0 (auto), -1 (none) or 1..4");
        ctx->opt_nsw = value;
    } else if (k == "stage_kb") {
        if (value < 0 || value > 64) return fail(QE_ERR_INVALID_ARGUMENT, "stage_kb must be in [0, 64] (0 = auto)");
        ctx->opt_stage_kb = value;
    } else if (k == "debug_flags") {
        ctx->opt_debug_flags = value;                       // timing experiments only (results are wrong)
    } else if (k == "stages") {
        if (value != 0 && (value < 2 || value > 32)) return fail(QE_ERR_INVALID_ARGUMENT, "stages must be 0 (auto) or 2..32");
        ctx->opt_stages = value;
    } else if (k == "jit_cache_entries") {
        if (value < 1 || value > 65536) return fail(QE_ERR_INVALID_ARGUMENT, "jit_cache_entries must be in [1, 65536]");
        ctx->opt_jit_cache_entries = value;
    } else return fail(QE_ERR_INVALID_ARGUMENT, "unknown option '" + k + "'");
    return QE_OK;
}

extern "C" int64_t qe_ctx_launch_count(const qe_ctx *ctx) { return ctx ? ctx->launches : 0; }

extern "C" int qe_ctx_trim(qe_ctx *ctx)
{
    if (!ctx) return fail(QE_ERR_INVALID_ARGUMENT, "ctx is NULL");
    DeviceGuard g(ctx->device);
    QE_CUDA(cudaStreamSynchronize(ctx->stream));             // freed blocks may still be in use by enqueued work
    ctx->pool.release_cached();
    return QE_OK;
}

extern "C" int qe_yardstick(qe_ctx *ctx, int kind, int64_t bytes, int reps, float *best_ms)
{
    if (!ctx || !best_ms) return fail(QE_ERR_INVALID_ARGUMENT, "NULL argument");
    if (kind < 0 || kind > 2 || bytes < (1 << 20) || reps < 1) return fail(QE_ERR_INVALID_ARGUMENT, "bad yardstick arguments");
    DeviceGuard g(ctx->device);
    bytes = bytes / 4096 * 4096;
    void *src = nullptr, *dst = nullptr, *sink = nullptr;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    int rc = QE_OK;
    float best = 1e30f;
    do {
        cudaError_t e = cudaSuccess;
        if (kind != 2 && (e = ctx->pool.alloc((size_t)bytes, &src)) != cudaSuccess) { rc = fail_cuda("yardstick allocation", e); break; }
        if (kind != 0 && (e = ctx->pool.alloc((size_t)bytes, &dst)) != cudaSuccess) { rc = fail_cuda("yardstick allocation", e); break; }
        if ((e = ctx->pool.alloc(64, &sink)) != cudaSuccess) { rc = fail_cuda("yardstick allocation", e); break; }
        if (src) cudaMemsetAsync(src, 1, (size_t)bytes, ctx->stream);
        cudaEventCreate(&e0);
        cudaEventCreate(&e1);
        for (int i = 0; i < reps + 1; i++) {                  // the first pass is a warm-up
            cudaEventRecord(e0, ctx->stream);
            if ((e = qe::launch_yardstick(src, dst, bytes, kind, ctx->num_sms, sink, ctx->stream)) != cudaSuccess) break;
            cudaEventRecord(e1, ctx->stream);
            if ((e = cudaEventSynchronize(e1)) != cudaSuccess) break;
            float ms = 0.f;
            cudaEventElapsedTime(&ms, e0, e1);
            if (i > 0) best = std::min(best, ms);
            ctx->launches++;
        }
        if (e != cudaSuccess) rc = fail_cuda("yardstick kernel", e);
    } while (false);
    if (e0) cudaEventDestroy(e0);
    if (e1) cudaEventDestroy(e1);
    cudaStreamSynchronize(ctx->stream);
    ctx->pool.free(src);
    ctx->pool.free(dst);
    ctx->pool.free(sink);
    *best_ms = best;
    return rc;
}

extern "C" int qe_jit_stats(double out[6])
{
    if (!out) return fail(QE_ERR_INVALID_ARGUMENT, "out is NULL");
    const qe::CubinStats st = qe::cubin_stats();
    for (int i = 0; i < 4; i++) out[i] = (double)st.hits[i];
    out[4] = st.compile_ms;
    out[5] = (double)qe::kAotCount;
    return QE_OK;
}

extern "C" int qe_host_alloc(size_t bytes, void **out)
{
    if (!out) return fail(QE_ERR_INVALID_ARGUMENT, "out is NULL");
    QE_CUDA(cudaHostAlloc(out, bytes ? bytes : 1, cudaHostAllocDefault));
    return QE_OK;
}

extern "C" int qe_host_free(void *p)
{
    if (p) QE_CUDA(cudaFreeHost(p));
    return QE_OK;
}

// ----------------------------------------------------------------- program

extern "C" int qe_program_create(qe_ctx *ctx, const uint32_t *words, size_t nwords, qe_program **out)
{
    if (!out) return fail(QE_ERR_INVALID_ARGUMENT, "out is NULL");
    *out = nullptr;
    if (!ctx) return fail(QE_ERR_INVALID_ARGUMENT, "ctx is NULL");
    qe_program *p = new qe_program();
    p->ctx = ctx;
    std::string err;
    if (!qe::parse_program(words, nwords, p->prog, err)) {
        delete p;
        return fail(QE_ERR_PROGRAM, err);
    }
    *out = p;
    return QE_OK;
}

extern "C" int qe_program_destroy(qe_program *program)
{
    delete program;
    return QE_OK;
}

// ------------------------------------------------------------------- table

extern "C" int qe_table_create(qe_ctx *ctx, int ncols, const qe_col_desc *cols, int64_t nrows, qe_table **out)
{
    if (!out) return fail(QE_ERR_INVALID_ARGUMENT, "out is NULL");
    *out = nullptr;
    if (!ctx || !cols || ncols <= 0) return fail(QE_ERR_INVALID_ARGUMENT, "ctx/cols is NULL or ncols <= 0");
    if (nrows < 0 || nrows > 0xffffffffLL)
        return fail(QE_ERR_INVALID_ARGUMENT, "nrows must fit the 32-bit selection vector (shard larger tables by row range)");
    for (int c = 0; c < ncols; c++) {
        if (cols[c].type < 0 || cols[c].type > 4) return fail(QE_ERR_INVALID_ARGUMENT, "unknown column type");
    }
    DeviceGuard g(ctx->device);
    qe_table *t = new qe_table();
    t->ctx = ctx;
    t->nrows = nrows;
    t->padded_rows = (nrows + kPadRows - 1) / kPadRows * kPadRows + kPadRows;
    t->cols.assign(cols, cols + ncols);
    t->values.assign(ncols, nullptr);
    t->valid.assign(ncols, nullptr);
    for (int c = 0; c < ncols; c++) {
        const size_t bytes = column_bytes(cols[c].type, t->padded_rows);
        cudaError_t e = ctx->pool.alloc(bytes, &t->values[c]);
        if (e == cudaSuccess && cols[c].nullable) {
            void *v = nullptr;
            e = ctx->pool.alloc(column_bytes(QE_TYPE_BOOLEAN, t->padded_rows), &v);
            t->valid[c] = (uint32_t *)v;
        }
        if (e != cudaSuccess) {
            (void)cudaGetLastError();
            qe_table_destroy(t);
            return fail_cuda("device allocation for table column", e);
        }
        // zero the padding (whole bitmaps; the tail of value arrays)
        if (cols[c].type == QE_TYPE_BOOLEAN) {
            cudaMemsetAsync(t->values[c], 0, bytes, ctx->stream);
        } else {
            const size_t w = type_width(cols[c].type);
            const int64_t tail0 = nrows / 128 * 128;
            cudaMemsetAsync((char *)t->values[c] + (size_t)tail0 * w, 0, (size_t)(t->padded_rows - tail0) * w, ctx->stream);
        }
        if (t->valid[c]) cudaMemsetAsync(t->valid[c], 0, column_bytes(QE_TYPE_BOOLEAN, t->padded_rows), ctx->stream);
    }
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { qe_table_destroy(t); return fail_cuda("cudaMemsetAsync", e); }
    *out = t;
    return QE_OK;
}

extern "C" int qe_table_destroy(qe_table *table)
{
    if (!table) return QE_OK;
    DeviceGuard g(table->ctx->device);
    cudaStreamSynchronize(table->ctx->stream);
    for (void *p : table->values) table->ctx->pool.free(p);
    for (uint32_t *p : table->valid) table->ctx->pool.free(p);
    delete table;
    return QE_OK;
}

static bool is_pinned(const void *p)
{
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { (void)cudaGetLastError(); return false; }
    return a.type == cudaMemoryTypeHost;
}

// host -> HBM.  Pinned sources go straight to cudaMemcpyAsync; pageable ones are
// staged through two pinned buffers so the copy engine always has a chunk in
// flight while the CPU fills the other.
static int upload_bytes(qe_ctx *ctx, void *dst, const void *src, size_t bytes)
{
    if (bytes == 0) return QE_OK;
    if (is_pinned(src)) {
        QE_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, ctx->stream));
        return QE_OK;
    }
    for (int i = 0; i < 2; i++) {
        if (!ctx->staging[i]) {
            QE_CUDA(cudaHostAlloc(&ctx->staging[i], kStagingBytes, cudaHostAllocDefault));
            QE_CUDA(cudaEventCreateWithFlags(&ctx->staging_free[i], cudaEventDisableTiming));
        }
    }
    size_t off = 0;
    int slot = 0;
    while (off < bytes) {
        const size_t n = std::min(kStagingBytes, bytes - off);
        QE_CUDA(cudaEventSynchronize(ctx->staging_free[slot]));       // previous copy out of this buffer done
        memcpy(ctx->staging[slot], (const char *)src + off, n);
        QE_CUDA(cudaMemcpyAsync((char *)dst + off, ctx->staging[slot], n, cudaMemcpyHostToDevice, ctx->stream));
        QE_CUDA(cudaEventRecord(ctx->staging_free[slot], ctx->stream));
        off += n;
        slot ^= 1;
    }
    return QE_OK;
}

extern "C" int qe_table_upload(qe_table *table, int col, const void *values, const uint8_t *validity,
                               int64_t row0, int64_t n)
{
    if (!table) return fail(QE_ERR_INVALID_ARGUMENT, "table is NULL");
    if (col < 0 || col >= (int)table->cols.size()) return fail(QE_ERR_INVALID_ARGUMENT, "column index out of range");
    if (row0 < 0 || n < 0 || row0 + n > table->nrows) return fail(QE_ERR_INVALID_ARGUMENT, "row range outside the table");
    const qe_col_desc &cd = table->cols[col];
    const bool bitmap_involved = cd.type == QE_TYPE_BOOLEAN || cd.nullable;
    if (bitmap_involved && (row0 % 8) != 0) return fail(QE_ERR_INVALID_ARGUMENT, "row0 must be a multiple of 8 for bitmap columns");
    qe_ctx *ctx = table->ctx;
    DeviceGuard g(ctx->device);
    int rc = QE_OK;
    if (values) {
        if (cd.type == QE_TYPE_BOOLEAN)
            rc = upload_bytes(ctx, (char *)table->values[col] + row0 / 8, values, (size_t)((n + 7) / 8));
        else
            rc = upload_bytes(ctx, (char *)table->values[col] + (size_t)row0 * type_width(cd.type), values,
                              (size_t)n * type_width(cd.type));
        if (rc) return rc;
    }
    if (cd.nullable) {
        char *dst = (char *)table->valid[col] + row0 / 8;
        const size_t nbytes = (size_t)((n + 7) / 8);
        if (validity) {
            rc = upload_bytes(ctx, dst, validity, nbytes);
        } else if (n > 0) {
            // all non-null: whole bytes of ones, then the partial last byte
            QE_CUDA(cudaMemsetAsync(dst, 0xff, (size_t)(n / 8), ctx->stream));
            if (n % 8) QE_CUDA(cudaMemsetAsync(dst + n / 8, (1 << (n % 8)) - 1, 1, ctx->stream));
        }
        if (rc) return rc;
    } else if (validity) {
        return fail(QE_ERR_INVALID_ARGUMENT, "validity bitmap given for a column declared non-nullable");
    }
    return QE_OK;
}

extern "C" void *qe_table_device_ptr(qe_table *table, int col, int which)
{
    if (!table || col < 0 || col >= (int)table->cols.size()) return nullptr;
    return which == 0 ? table->values[col] : (void *)table->valid[col];
}

extern "C" int qe_table_generate(qe_table *table, int col, int kind, uint64_t seed, double p0, double p1,
                                 double null_rate, int64_t row_offset)
{
    if (!table) return fail(QE_ERR_INVALID_ARGUMENT, "table is NULL");
    if (col < 0 || col >= (int)table->cols.size()) return fail(QE_ERR_INVALID_ARGUMENT, "column index out of range");
    if (kind < 0 || kind > 4) return fail(QE_ERR_INVALID_ARGUMENT, "unknown generator kind");
    // INT64 columns widen exactly to the reference's DOUBLE only below 2^53 (include/qegpu.h, qe_type)
    if (table->cols[col].type == QE_TYPE_INT64 && kind == QE_GEN_UNIFORM_INT &&
        (std::fabs(p0) >= 9007199254740992.0 || std::fabs(p1) >= 9007199254740992.0))
        return fail(QE_ERR_INVALID_ARGUMENT, "INT64 columns must hold |x| < 2^53 (they stand for the reference's DOUBLE values)");
    qe_ctx *ctx = table->ctx;
    DeviceGuard g(ctx->device);
    QE_CUDA(qe::launch_generate(table->values[col], table->valid[col], table->cols[col].type, kind, seed, p0, p1,
                                table->cols[col].nullable ? null_rate : 0.0, row_offset, table->nrows, ctx->stream));
    ctx->launches++;
    return QE_OK;
}

// ------------------------------------------------------------------ result

struct DevBuf {
    void *ptr = nullptr;
    size_t bytes = 0;
};

struct ResultColumn {
    int type = 0;
    bool has_validity = false;
    DevBuf values, valid;
};

// host-visible words the kernels' results are copied into (pinned)
struct ResultHost {
    unsigned long long count;
    double agg[QE_MAX_AGGS][4];
};

struct qe_result {
    qe_ctx *ctx = nullptr;
    int mode = 0;                       // qe::KernelMode of the last run
    bool used_jit = false;
    int64_t batch_rows = 0;
    int64_t capacity = 0;
    std::string f_sig;                  // filter runs: signature of the last pipeline run into this result ...
    double f_sel = -1.0;                // ... and the selectivity it measured (kept / rows; -1 unknown)
    std::vector<ResultColumn> cols;
    std::vector<int> agg_fns;
    DevBuf sel, ctl, partials, final_aggs, debug;
    ResultHost *host = nullptr;         // pinned
    cudaEvent_t ev_begin = nullptr, ev_end = nullptr, ev_done = nullptr;
    bool pending = false;
    bool chained = false;               // a chunk of a bounded qe_pipeline_stream: host->count is the running total of the stream
    bool streamed = false;              // outputs went to the caller's host buffers chunk by chunk; none are held on the device
    // aggregation: super-tile partials [n_super][aggregates][4] (pinned), folded in row order on first access
    double *h_super = nullptr;
    size_t h_super_cap = 0;             // doubles
    int64_t n_super = 0;
    bool agg_folded = true;
    std::vector<qe_result *> ring;      // bounded streaming: further output buffer sets, owned by this result
    bool timed = false;                 // ev_begin / ev_end hold a run not yet folded into the totals
    double total_kernel_ms = 0.0;       // device time of the main kernel over all runs into this result
    int64_t kernel_runs = 0;
    // group-by: the table of groups on the device, and the finished rows (one per group) on the host
    DevBuf g_state, g_keys, g_nulls, g_acc, g_ctl, g_part, g_part_slot, g_records;
    DevBuf g_pt_words, g_pt_cursor;                          // two-pass variant: the partitioned records
    std::vector<std::vector<uint8_t>> h_values, h_valid;    // per output column: values in the column's layout, Arrow bitmap
    int64_t h_rows = 0;
    int g_copies = 0;                                        // shared-memory variants: copies of the accumulator words per CTA
    int g_nkeys = 0, g_variant = 0, g_threads = 0;          // variant: 0 private accumulators, 1 global atomics, 2 CTA-shared, 3 registers, 4 lane copies, 5 eight copies
    std::string g_seen_sig;                                 // signature of the last group-by that completed into this result ...
    int g_seen_groups = 0;                                  // ... and how many groups it found (sizes the register variant next time)
    std::string g_overflowed;                               // signature of the group-by whose private variant overflowed here
    int g_level = 0;                                        // ... and the variant to start with next time (0: private 512 threads,
                                                            // 1: private 256, 2: CTA-shared accumulators, 3: global atomics)
    std::vector<qe_agg_partial> g_partials;                 // [group][aggregate]
    std::vector<int64_t> g_first_row;                       // [group] table row of the group's first appearance
};

static bool is_groupby(const qe_result *r)
{
    return r->mode == qe::KM_GROUPBY_PRIVATE || r->mode == qe::KM_GROUPBY_ATOMIC || r->mode == qe::KM_GROUPBY_SHARED ||
           r->mode == qe::KM_GROUPBY_REG;
}

// Fold the last run's kernel time into the totals (the run must have completed).
static void fold_timing(qe_result *r)
{
    if (!r->timed) return;
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, r->ev_begin, r->ev_end) == cudaSuccess) {
        r->total_kernel_ms += ms;
        r->kernel_runs += 1;
    } else {
        (void)cudaGetLastError();
    }
    r->timed = false;
}

static int ensure(qe_ctx *ctx, DevBuf &b, size_t bytes)
{
    if (b.ptr && b.bytes >= bytes) return QE_OK;
    if (b.ptr) ctx->pool.free(b.ptr);
    b.ptr = nullptr;
    b.bytes = 0;
    cudaError_t e = ctx->pool.alloc(bytes, &b.ptr);
    if (e != cudaSuccess) { (void)cudaGetLastError(); return fail_cuda("device allocation for result", e); }
    b.bytes = bytes;
    return QE_OK;
}

static void release(qe_ctx *ctx, DevBuf &b)
{
    if (b.ptr) ctx->pool.free(b.ptr);
    b.ptr = nullptr;
    b.bytes = 0;
}

static int ensure_host_super(qe_result *r, size_t doubles)
{
    if (r->h_super && r->h_super_cap >= doubles) return QE_OK;
    if (r->h_super) {
        if (r->pending) { QE_CUDA(cudaEventSynchronize(r->ev_done)); r->pending = false; }   // a copy into it may be in flight
        cudaFreeHost(r->h_super);
    }
    r->h_super = nullptr;
    r->h_super_cap = 0;
    const size_t cap = std::max<size_t>(doubles, 1024);
    QE_CUDA(cudaHostAlloc((void **)&r->h_super, cap * sizeof(double), cudaHostAllocDefault));
    memset(r->h_super, 0, cap * sizeof(double));              // count 0 = no contribution
    r->h_super_cap = cap;
    return QE_OK;
}

// Third level of the fixed-shape aggregation: the super-tile partials in ascending row order, one addition each
// (the reference adds row by row in this order, M/evaluator/Accumulators.kt:44; here the rows of a super-tile arrive
// pre-added in a fixed tree).  qe_agg_combine over the concatenated lists of several partitions does the same.
static void fold_aggs(qe_result *r)
{
    if (r->mode != qe::KM_AGG || r->agg_folded) return;
    const size_t nagg = r->agg_fns.size();
    for (size_t k = 0; k < nagg && k < QE_MAX_AGGS; k++) {
        double sum = 0.0, mn = qe_bits_f64((qe_i64)0x7ff0000000000000LL), mx = qe_bits_f64((qe_i64)0xfff0000000000000ULL);
        int64_t count = 0;
        for (int64_t s = 0; s < r->n_super; s++) {
            const double *src = r->h_super + ((size_t)s * nagg + k) * 4;
            int64_t cnt;
            memcpy(&cnt, &src[1], sizeof cnt);
            if (cnt == 0) continue;
            count += cnt;
            sum += src[0];
            mn = qe_math_min(mn, src[2]);
            mx = qe_math_max(mx, src[3]);
        }
        r->host->agg[k][0] = sum;
        memcpy(&r->host->agg[k][1], &count, sizeof count);
        r->host->agg[k][2] = mn;
        r->host->agg[k][3] = mx;
    }
    r->agg_folded = true;
}

extern "C" int qe_result_destroy(qe_result *r)
{
    if (!r) return QE_OK;
    DeviceGuard g(r->ctx->device);
    if (r->pending) cudaEventSynchronize(r->ev_done);
    for (qe_result *m : r->ring) qe_result_destroy(m);
    r->ring.clear();
    if (r->h_super) cudaFreeHost(r->h_super);
    for (ResultColumn &c : r->cols) { release(r->ctx, c.values); release(r->ctx, c.valid); }
    release(r->ctx, r->sel);
    release(r->ctx, r->ctl);
    for (DevBuf *b : {&r->g_state, &r->g_keys, &r->g_nulls, &r->g_acc, &r->g_ctl, &r->g_part, &r->g_part_slot, &r->g_records, &r->g_pt_words, &r->g_pt_cursor})
        release(r->ctx, *b);
    release(r->ctx, r->partials);
    release(r->ctx, r->final_aggs);
    release(r->ctx, r->debug);
    if (r->host) cudaFreeHost(r->host);
    if (r->ev_begin) cudaEventDestroy(r->ev_begin);
    if (r->ev_end) cudaEventDestroy(r->ev_end);
    if (r->ev_done) cudaEventDestroy(r->ev_done);
    delete r;
    return QE_OK;
}

// ---------------------------------------------------------------- pipeline

static qe::TileShape choose_shape(const qe_ctx *ctx, const qe::BoundPipeline &bp, qe::KernelMode mode, double sel_hint = -1.0)
{
    // Aim for ~128 bytes of column data in flight per thread: enough 16/32-byte
    // requests outstanding to cover HBM latency at 8+ warps per scheduler,
    // without pushing the register count past 64 (4 CTAs of 256 threads per SM).
    size_t row_bytes = 0;
    for (const qe::ColumnInfo &c : bp.slot_info) row_bytes += c.type == QE_TYPE_BOOLEAN ? 1 : type_width(c.type);
    if (row_bytes == 0) row_bytes = 8;
    qe::TileShape s;
    size_t rows_per_thread = std::min<size_t>(std::max<size_t>(128 / row_bytes, 2), 32);
    if (mode == qe::KM_FILTER) rows_per_thread = std::max<size_t>(rows_per_thread, 8);   // amortise scan + lookback
    // group-by, private variant: 8 warps per SM only (shared memory holds the accumulators), so more rows per
    // thread in flight (measured: 48-byte rows 8 rows, 16-byte rows 4 rows per thread)
    if (mode == qe::KM_GROUPBY_PRIVATE) rows_per_thread = row_bytes >= 32 ? 8 : 4;
    size_t pow2 = 2;
    while (pow2 * 2 <= rows_per_thread) pow2 *= 2;
    s.vec = pow2 >= 4 ? 4 : 2;
    s.steps = (int)(pow2 / (size_t)s.vec);
    if (ctx->opt_vec) s.vec = (int)ctx->opt_vec;
    if (ctx->opt_steps) s.steps = (int)ctx->opt_steps;
    if (s.vec == 4 && !qe::jit_wide_loads_supported()) {      // older NVRTC: same rows per thread with 128-bit accesses
        s.vec = 2;
        if (!ctx->opt_steps) s.steps = std::min(8, s.steps * 2);
    }
    // Filter kernel: land the columns in shared-memory stages with bulk copies, so the bytes in
    // flight are not bounded by registers and the next tiles load during scan / look-back / scatter.
    s.staged = mode == qe::KM_FILTER && ctx->opt_staging == 1;
    s.pipelined = ctx->opt_pipeline != 0;
    if (s.staged) {
        s.vec = 2;                                             // staged rows are read back with 128-bit LDS
        if (!ctx->opt_steps) {                                 // ~16 KB per stage, 8 rows per thread at most
            const size_t rows = std::max<size_t>(512, std::min<size_t>(2048, (16u << 10) / row_bytes / 512 * 512));
            s.steps = (int)(rows / 512);
        }
        const size_t stage = qe::staged_smem_bytes(bp, s) / (size_t)s.nstages;
        size_t n = ctx->opt_stages ? (size_t)ctx->opt_stages : std::min<size_t>(4, std::max<size_t>(2, (48u << 10) / stage));
        while (n > 2 && n * stage > (200u << 10)) n--;
        s.nstages = (int)n;
        if ((size_t)s.nstages * stage > (220u << 10)) s.staged = false;   // does not fit: register-sourced kernel
    }
    s.min_blocks = ctx->opt_min_blocks ? (int)ctx->opt_min_blocks : 1;
    // Filter kernel, default: persistent warp-specialised CTA around a ring of bulk-copy stages.
    s.ring = mode == qe::KM_FILTER && (ctx->opt_staging == -1 || ctx->opt_staging == 2);
    if (s.ring) {
        // consumer warps: 16, or 20 for a single 8-byte column -- at low selectivity that kernel is bound by
        // issue slots and latency, not by HBM (1 %: 1.45 ms with 16 warps, 1.31 with 20 or 24; 99 %: 2.98 / 3.00 /
        // 3.02); 40-byte rows lose with more warps (6.9 ms with 16, 7.7 with 24): profiles/r01_sweep_06_*.jsonl
        // ... unless the same pipeline kept more than 40 % of its rows the last time it ran into this result: from there
        // on the kernel is bound by HBM and 16 warps are the faster shape (99 %: 2.98 against 3.00-3.04 ms)
        // A query that is known to keep little (at most 8 % the last time) runs on the row lists and the sweepers below: its
        // consumer warps pay a fixed price per tile (barriers, counts, the look-back behind them) that is as large as the
        // work of a 7-step slice, so larger tiles, more consumer warps and a third sweeper warp win there.  24 consumer
        // warps, 5 376-row / 42 KB tiles (five stages), three sweepers: 0.5 / 1 / 3 / 5 / 8 % 1.14-1.16 / 1.178 / 1.25 / 1.30 /
        // 1.40 ms per 10^9 rows, against 1.238 / 1.253 / 1.33 / 1.38 / 1.44 ms with the default shape (20 warps, 4 480-row
        // tiles, two sweepers) and 1.175 / 1.20 / 1.30 / 1.35 with 20 warps and 5 120-row tiles; at 10 % the default is as fast
        // (1.49-1.50) and from there on the smaller tiles win (50 %: 2.14 against 2.53 ms) --
        // profiles/r02_sweep_07_tile_rows.jsonl
        const bool known_sparse = row_bytes <= 8 && sel_hint >= 0.0 && sel_hint <= 0.08 && ctx->opt_list_cap >= 0 && ctx->opt_nsw >= 0;
        s.ncw = ctx->opt_ncw ? (int)ctx->opt_ncw : known_sparse ? 24 : ((row_bytes <= 8 && !(sel_hint > 0.4)) ? 20 : 16);
        const size_t quantum = (size_t)std::max(128, 32 * s.ncw);       // rows: whole warp steps, 16-byte bitmap runs
        const size_t stage_target = (size_t)(ctx->opt_stage_kb ? ctx->opt_stage_kb : 32) << 10;
        // One look-back per tile costs ~1500 cycles whatever the tile holds, so a tile must carry >= ~32 KB for
        // the CTA to keep its share of HBM busy: round UP to whole quanta while 5 stages still fit (40-byte rows:
        // 1024 rows = 40 KB measured 96.7 % of the copy peak, 512 rows = 20 KB 64.5 %), down otherwise.
        size_t rows = stage_target / row_bytes / quantum * quantum;
        if (rows * row_bytes < stage_target && (rows + quantum) * row_bytes <= (44u << 10)) rows += quantum;
        if (known_sparse && !ctx->opt_stage_kb) rows = std::max(rows, (size_t)(42u << 10) / row_bytes / quantum * quantum);
        if (ctx->opt_tile_rows) rows = (size_t)ctx->opt_tile_rows;
        rows = std::min<size_t>(std::max(rows, quantum), (size_t)kPadRows / quantum * quantum);
        if (rows % quantum) rows = (rows + quantum - 1) / quantum * quantum;
        s.tile_rows = (int)rows;
        s.nstages = 1;
        const size_t stage = qe::staged_smem_bytes(bp, s);
        // sparse tiles: packed row lists instead of keep masks (qe_ring_scatter_list, sweeper warps)
        bool bit_outputs = false;
        for (const qe::BoundOutput &o : bp.outs) bit_outputs |= o.type == QE_TYPE_BOOLEAN || o.has_validity;
        // Measured on B200, one 8-byte column, 10^9 rows (profiles/r02_sweep_02_sweepers.jsonl): 1 / 2 / 5 / 10 % kept
        // 1.32 / 1.50 / 1.61 / 1.64 ms with the masks only, 1.28 / 1.31 / 1.40 / 1.58 ms with lists of 32 and two sweeper
        // warps; from ~12 % up most warps keep more than 32 rows of a tile and the mixed kernel loses (15 %: 1.79 against
        // 1.65 ms; lists of 64 lose everywhere above 5 %).  Wide rows are bound by HBM, not by the consumers' instruction
        // issue: the 40-byte lineitem filter (24 % kept) runs at the copy peak with masks, 3.46 ms per 5*10^8 rows, and at
        // 3.59 ms with lists and sweepers (profiles/r02_sweep_03_lineitem_lists.jsonl).  So: lists for rows of up to 16
        // bytes, unless the same query kept more than 13.5 % last time (with the later list-building and sweeper changes the
        // lists win up to 13 %: 1.60 against 1.65 ms; 15 %: 1.71 against 1.65).
        s.list_cap = ctx->opt_list_cap < 0 ? 0 : ctx->opt_list_cap ? (int)ctx->opt_list_cap : ((row_bytes > 16 || sel_hint > 0.135) ? 0 : 32);
        if (bit_outputs && ctx->opt_nsw < 0) s.list_cap = 0;              // compacted bit columns of listed rows are written by the sweepers only
        s.nsw = !s.list_cap || ctx->opt_nsw < 0 ? 0 : ctx->opt_nsw ? (int)ctx->opt_nsw : (known_sparse ? 3 : 2);
        const size_t entry = (size_t)s.tile_rows / (size_t)s.ncw <= 256 ? 1 : 2;                                   // bytes per list entry
        const size_t meta = 64 + (size_t)s.tile_rows / 8 + (4 + 8 + entry * (size_t)s.list_cap) * (size_t)s.ncw;   // per stage: barriers, keep masks, counts, bases, lists
        size_t n = ctx->opt_stages ? (size_t)ctx->opt_stages : 16;
        while (n > 2 && n * (stage + meta) > (224u << 10)) n--;
        s.nstages = (int)n;
        s.defer = ctx->opt_defer ? (int)ctx->opt_defer : (n >= 6 ? 3 : 2);   // leave >= 3 stages to the loads in flight
        if (s.defer >= s.nstages) s.defer = s.nstages - 1;
        s.loads = (int)ctx->opt_loads;
        s.nlb = std::min(ctx->opt_nlb ? (int)ctx->opt_nlb : 1, s.nstages);   // the end markers need a stage each
        while (2 + s.nlb + s.nsw + s.ncw > 32 && s.nsw > 1) s.nsw--;         // a CTA has at most 32 warps
        while (2 + s.nlb + s.nsw + s.ncw > 32 && s.nlb > 1) s.nlb--;
        if (n * (stage + meta) > (224u << 10) || s.defer < 1) s.ring = false;   // a single tile does not fit: register-sourced kernel
    }
    // Dense projection through the same ring of stages (one producer warp + 16 consumer warps): on request only
    // ("staging" = 2).  Measured equal to the grid-stride kernel on the nullable tree (6.54 against 6.51 ms per
    // 10^9 rows, both at the DRAM rate of an 80 % read / 20 % write mix), so the simpler kernel stays the default.
    if (mode == qe::KM_PROJECT && ctx->opt_staging == 2 && !ctx->opt_steps && !ctx->opt_vec) {
        s.pring = true;
        s.ncw = ctx->opt_ncw ? (int)ctx->opt_ncw : 16;
        const size_t quantum = (size_t)std::max(128, 32 * s.ncw);
        const size_t stage_target = (size_t)(ctx->opt_stage_kb ? ctx->opt_stage_kb : 32) << 10;
        size_t rows = stage_target / row_bytes / quantum * quantum;
        if (rows * row_bytes < stage_target && (rows + quantum) * row_bytes <= (44u << 10)) rows += quantum;
        if (ctx->opt_tile_rows) rows = (size_t)ctx->opt_tile_rows;
        rows = std::min<size_t>(std::max(rows, quantum), (size_t)kPadRows / quantum * quantum);
        if (rows % quantum) rows = (rows + quantum - 1) / quantum * quantum;
        s.tile_rows = (int)rows;
        s.nstages = 1;
        const size_t stage = qe::staged_smem_bytes(bp, s);
        size_t n = ctx->opt_stages ? (size_t)ctx->opt_stages : 8;
        while (n > 2 && n * stage > (216u << 10)) n--;
        s.nstages = (int)n;
        if (n * stage > (216u << 10) || s.ncw % 4 || bp.slot_info.empty()) s.pring = false;   // does not fit (or reads no column): grid-stride kernel
    }
    // Group-by, register variant: the same ring of stages, one producer warp + ncw consumer warps, no look-back.
    if (mode == qe::KM_GROUPBY_REG) {
        s.greg = true;
        // Registers are allocated per four warps: 1 + 11 warps get 168 registers per thread, 13-16 warps 128 (which
        // spills the accumulators).  Eleven consumer warps need tiles of 1408 rows (whole warp steps and 16-byte
        // runs of every column and bitmap); when three such stages do not fit, eight warps and 256-row quanta.
        s.ncw = ctx->opt_ncw ? (int)ctx->opt_ncw : 11;
        if (s.ncw == 11 && 3 * 1408 * row_bytes > (200u << 10)) s.ncw = 8;
        const size_t quantum = s.ncw == 11 ? 1408 : (size_t)std::max(128, 32 * s.ncw);
        const size_t stage_target = (size_t)(ctx->opt_stage_kb ? ctx->opt_stage_kb : 32) << 10;
        size_t rows = stage_target / row_bytes / quantum * quantum;
        if (rows * row_bytes < stage_target && (rows + quantum) * row_bytes <= (44u << 10)) rows += quantum;
        if (ctx->opt_tile_rows) rows = (size_t)ctx->opt_tile_rows;
        rows = std::min<size_t>(std::max(rows, quantum), (size_t)kPadRows / quantum * quantum);
        if (rows % quantum) rows = (rows + quantum - 1) / quantum * quantum;
        s.tile_rows = (int)rows;
        s.nstages = 1;
        const size_t stage = qe::staged_smem_bytes(bp, s);
        size_t n = ctx->opt_stages ? (size_t)ctx->opt_stages : 8;
        while (n > 2 && n * stage > (208u << 10)) n--;         // the key cache and the barriers are static shared memory
        s.nstages = (int)n;
        if (n * stage > (208u << 10)) s.greg = false;
    }
    return s;
}

static void serialize_interp(const qe::BoundPipeline &bp, std::vector<uint32_t> &w)
{
    w.assign(QE_IP_PROGRAMS, 0);
    w[0] = bp.has_filter ? 1 : 0;
    w[1] = (uint32_t)bp.proj.size();
    w[2] = (uint32_t)bp.agg_fns.size();
    w[3] = (uint32_t)bp.slot_info.size();
    for (size_t s = 0; s < bp.slot_info.size(); s++)
        w[QE_IP_SLOTS + s] = (uint32_t)bp.slot_info[s].type | ((bp.slot_info[s].nullable ? 1u : 0u) << 8);
    for (size_t k = 0; k < bp.proj.size(); k++) {
        w[QE_IP_OUTS + k] = (uint32_t)bp.outs[k].type | ((bp.outs[k].has_validity ? 1u : 0u) << 8) |
                            ((uint32_t)(bp.proj[k].identity_slot + 1) << 16);
        if (k < bp.agg_fns.size()) w[QE_IP_AGGS + k] = (uint32_t)bp.agg_fns[k];
    }
    auto put = [&](const qe::BoundExpr &e) {
        const size_t at = w.size();
        w.push_back(0);
        for (const qe::Instr &in : e.code) {
            w.push_back((uint32_t)in.op | (in.arg << 8));
            if (in.op == qe::OP_CONST_F64) {
                uint64_t bits;
                memcpy(&bits, &in.k, sizeof bits);
                w.push_back((uint32_t)bits);
                w.push_back((uint32_t)(bits >> 32));
            }
            if (in.op == qe::OP_CONST_STR) w.push_back((uint32_t)in.code);
        }
        w[at] = (uint32_t)(w.size() - at - 1);
    };
    if (bp.has_filter) put(bp.filter);
    for (const qe::BoundExpr &e : bp.proj) put(e);
}

// Generated kernel for (pipeline, mode, tile shape) on this context's device.  Returns QE_OK, QE_ERR_JIT when the
// compilation or the load failed, or -1 when the cubin is in no cache and there is no NVRTC to make it (the caller
// then runs the interpreter kernels).
static int get_jit_kernel(qe_ctx *ctx, const qe::BoundPipeline &bp, qe::KernelMode mode, const qe::TileShape &shape,
                          JitEntry **out, size_t dyn_smem = 0)
{
    char suffix[192];
    snprintf(suffix, sizeof suffix, "|M%d|B%d|S%d|V%d|T%d.%d|D%d|P%d|R%d.%d.%d.%d.%d.%d|G%d", (int)mode, shape.min_blocks, shape.steps, shape.vec, shape.staged ? 1 : 0,
             shape.nstages, (int)ctx->opt_debug_flags, shape.pipelined ? 1 : 0, shape.ring ? 1 : 0, shape.tile_rows, shape.ncw,
             shape.defer, shape.loads, shape.nlb, shape.gthreads + (shape.greg ? 100000 + 1000 * shape.rmax_hint : 0) + (shape.pring ? 200000 : 0) + (shape.scopies < 0 ? 1600000 : shape.scopies > 1 ? 400000 * shape.scopies + 100000000 * shape.lane_acc_kb : 0));
    snprintf(suffix + strlen(suffix), sizeof suffix - strlen(suffix), "|Q%d.%d.%d|L%d.%d", shape.gpass, shape.npart, shape.pcap, shape.ring ? shape.list_cap : 0,
             (shape.ring && shape.list_cap) ? shape.nsw : 0);
    const std::string key = bp.signature() + suffix;
    auto it = ctx->jit_cache.find(key);
    if (it != ctx->jit_cache.end()) {
        ctx->jit_lru.splice(ctx->jit_lru.end(), ctx->jit_lru, it->second.lru);      // most recently used
        *out = &it->second;
        return QE_OK;
    }
    const std::string src = qe::generate_source(bp, mode, shape);
    std::vector<std::string> extra;
    if (ctx->opt_debug_flags) extra.push_back("-DQE_DEBUG_FLAGS=" + std::to_string(ctx->opt_debug_flags));
    qe::CubinRef cubin;
    std::string log;
    int origin = 0;
    const int crc = qe::cubin_get(src, extra, cubin, &origin, log);
    if (crc < 0) { tls_error = "no cached kernel for this pipeline and no compiler: " + log; return -1; }
    if (crc) {
        if (getenv("QE_DUMP_JIT")) fprintf(stderr, "%s\n", src.c_str());
        return fail(QE_ERR_JIT, "compiling the generated pipeline failed: " + log);
    }
    // bounded: the least recently used kernels are unloaded (nothing of theirs may still be running)
    while ((int64_t)ctx->jit_cache.size() >= ctx->opt_jit_cache_entries && !ctx->jit_lru.empty()) {
        cudaStreamSynchronize(ctx->stream);
        auto victim = ctx->jit_cache.find(ctx->jit_lru.front());
        if (victim != ctx->jit_cache.end()) {
            qe::jit_unload(victim->second.kernel);
            ctx->jit_cache.erase(victim);
        }
        ctx->jit_lru.pop_front();
    }
    JitEntry e;
    e.shape = shape;
    std::string err;
    if (!qe::jit_load(*cubin, qe::jit_entry_name(mode, shape), dyn_smem ? dyn_smem : qe::staged_smem_bytes(bp, shape),
                      (mode == qe::KM_FILTER || mode == qe::KM_GROUPBY_REG || shape.pring) ? shape.block_threads()
                      : (mode == qe::KM_GROUPBY_PRIVATE || mode == qe::KM_GROUPBY_SHARED) ? shape.gthreads : QE_BLOCK,
                      e.kernel, err))
        return fail(QE_ERR_JIT, err);
    ctx->jit_loads++;
    if (getenv("QE_DUMP_JIT"))
        fprintf(stderr, "[qegpu] kernel %s: origin %d (0 memory, 1 aot, 2 disk, 3 nvrtc), %d regs, %d CTAs/SM\n", key.c_str(), origin,
                e.kernel.regs, e.kernel.blocks_per_sm);
    ctx->jit_lru.push_back(key);
    e.lru = std::prev(ctx->jit_lru.end());
    auto ins = ctx->jit_cache.emplace(key, e);
    *out = &ins.first->second;
    return QE_OK;
}

// literal values of the bound pipeline -> kernel parameters (see Literal in qe_program.h)
static void fill_literals(const qe::BoundPipeline &bp, QeParams &p)
{
    for (size_t i = 0; i < bp.lits.size() && i < QE_MAX_LITS; i++) {
        const qe::Literal &l = bp.lits[i];
        if (l.is_str) { p.lit_lo[i] = l.code; p.lit_hi[i] = l.code; continue; }
        p.lit_f64[i] = l.k;
        if (l.int_ok()) { p.lit_lo[i] = (long long)std::floor(l.k); p.lit_hi[i] = (long long)std::ceil(l.k); }
        if (l.mod_ok()) { p.lit_abs[i] = std::fabs(l.k); p.lit_inv[i] = 1.0 / std::fabs(l.k); }
    }
}

// signature + literal values: "the same query", for the hints a reused result carries from its last run
static std::string query_key(const qe::BoundPipeline &bp)
{
    std::string s = bp.signature() + "|L";
    char b[24];
    for (const qe::Literal &l : bp.lits) {
        unsigned long long bits;
        memcpy(&bits, &l.k, sizeof bits);
        snprintf(b, sizeof b, "%llx.%d,", bits, l.code);
        s += b;
    }
    return s;
}

// What qe_pipeline_stream adds to a run (not part of the ABI).
struct RunExtra {
    const qe_u64 *chain_base = nullptr;   // bounded streaming: device word holding the rows kept by the earlier chunks; this
                                          // run writes into buffers of its own and keeps only the bit phase (out_base_mask 7)
    bool chained = false;
    int64_t super_offset = 0;             // aggregation: where this chunk's super-tile partials go in the host list ...
    int64_t super_total = 0;              // ... and the length of the whole list (0: this run alone)
};

static int ensure_host_super(qe_result *r, size_t doubles);

static int run_pipeline(qe_ctx *ctx, const qe_table *table, const qe_pipeline_desc *d, qe_result **result, const RunExtra *extra)
{
    if (!ctx || !table || !d || !result) return fail(QE_ERR_INVALID_ARGUMENT, "NULL argument");
    if (table->ctx != ctx) return fail(QE_ERR_INVALID_ARGUMENT, "table belongs to another context");
    if (d->nproj < 0 || d->nproj > QE_MAX_PROJECTIONS) return fail(QE_ERR_UNSUPPORTED, "nproj out of range");
    if (d->nagg < 0 || (d->nagg > 0 && !d->agg_fns)) return fail(QE_ERR_INVALID_ARGUMENT, "bad aggregate list");
    DeviceGuard g(ctx->device);

    // ---- bind: resolve columns, type-check, decide output types
    std::vector<qe::ColumnInfo> schema(table->cols.size());
    for (size_t c = 0; c < table->cols.size(); c++) {
        schema[c].type = table->cols[c].type;
        schema[c].nullable = table->cols[c].nullable != 0;
    }
    // QE_RUN_NO_VALUES: the caller holds the table (MemoryTable lives in host memory, M/data/MemoryTable.kt:7) and
    // FilterOperator hands back the very row it was given (M/operator/FilterOperator.kt:20-22), so the selection
    // vector alone is its output: the projections are neither evaluated nor written
    const bool no_values = (d->flags & QE_RUN_NO_VALUES) && d->filter && d->nagg == 0;
    std::vector<const qe::Program *> proj;
    for (int k = 0; k < d->nproj && !no_values; k++) {
        if (!d->proj || !d->proj[k]) return fail(QE_ERR_INVALID_ARGUMENT, "NULL projection program");
        proj.push_back(&d->proj[k]->prog);
    }
    std::vector<int> aggs(d->agg_fns, d->agg_fns + d->nagg);
    qe::BoundPipeline bp;
    std::string err;
    int rc = qe::bind_pipeline(schema, d->filter ? &d->filter->prog : nullptr, proj, aggs, bp, err);
    if (rc) return fail(rc, err);

    const qe::KernelMode mode = !bp.agg_fns.empty() ? qe::KM_AGG : bp.has_filter ? qe::KM_FILTER : qe::KM_PROJECT;

    // ---- batch
    const int64_t row_begin = d->row_begin;
    const int64_t row_end = d->row_end < 0 ? table->nrows : d->row_end;
    if (row_begin < 0 || row_end < row_begin || row_end > table->nrows) return fail(QE_ERR_INVALID_ARGUMENT, "bad row range");
    if (row_begin % 128) return fail(QE_ERR_INVALID_ARGUMENT, "row_begin must be a multiple of 128");
    const int64_t n = row_end - row_begin;
    if (mode == qe::KM_FILTER && !(d->flags & QE_RUN_NO_SELECTION) &&
        (row_begin + d->sel_offset < 0 || row_end + d->sel_offset > 0x100000000LL))
        return fail(QE_ERR_INVALID_ARGUMENT, "selection vector values (row + sel_offset) must fit 32 bits");

    // ---- kernel: generated code (from the caches or NVRTC), or the interpreter
    bool use_jit = ctx->opt_jit != 0 && !(d->flags & QE_RUN_INTERPRETER);
    qe::TileShape shape;
    JitEntry *jit = nullptr;
    int blocks_per_sm = 1;
    const std::string qkey = mode == qe::KM_FILTER ? query_key(bp) : std::string();
    if (use_jit) {
        // a reused result knows the selectivity of its last run (waiting for that run is what a reuse does anyway)
        double sel_hint = -1.0;
        if (qe_result *prev = *result; prev && prev->ctx == ctx && mode == qe::KM_FILTER && prev->mode == (int)mode &&
                                       !(d->flags & QE_RUN_APPEND) && !(extra && extra->chained) && prev->f_sig == qkey) {
            if (prev->pending) { QE_CUDA(cudaEventSynchronize(prev->ev_done)); prev->pending = false; }
            if (prev->batch_rows > 0 && (int64_t)prev->host->count <= prev->capacity)
                prev->f_sel = (double)prev->host->count / (double)prev->batch_rows;
            sel_hint = prev->f_sel;
        } else if (qe_result *prev = *result; prev && prev->f_sig == qkey) {
            sel_hint = prev->f_sel;                           // appended / chained chunks keep the shape of the first
        }
        if (sel_hint < 0.0 && mode == qe::KM_FILTER && !(d->flags & QE_RUN_APPEND)) {
            auto memo = ctx->sel_memo.find(qkey);
            if (memo != ctx->sel_memo.end()) sel_hint = memo->second;
        }
        shape = choose_shape(ctx, bp, mode, sel_hint);
        // Dense projection: four CTAs per SM (64 registers) hide the latency of the FP64 divide / MOD sequences better
        // than three; kept unless the tighter bound spills more than a few words.  (Nullable tree, 1e9 rows: 84
        // registers / three CTAs 7.06 ms, 64 registers with 16 bytes of stack / four CTAs 6.14 ms, __launch_bounds__(256, 3)
        // 6.48 ms -- profiles/r02_tree_*.log.  With literals as kernel parameters the 64-register build keeps two values on
        // the stack; refusing it for that cost 13 %.)
        rc = QE_OK;
        if (mode == qe::KM_PROJECT && !ctx->opt_min_blocks && !shape.pring) {
            qe::TileShape tight = shape;
            tight.min_blocks = 4;
            JitEntry *cand = nullptr;
            rc = get_jit_kernel(ctx, bp, mode, tight, &cand);
            if (rc == QE_OK && cand->kernel.local_bytes <= 64) { shape = tight; jit = cand; }
            else if (rc > 0) return rc;
        }
        if (!jit && rc >= 0) rc = get_jit_kernel(ctx, bp, mode, shape, &jit);
        if (rc > 0) return rc;
        if (rc < 0) {                                          // no compiler and not in the AOT table / disk cache
            if (!ctx->warned_no_nvrtc) {
                fprintf(stderr, "[qegpu] warning: %s; using the interpreter kernels (slower, still on the GPU)\n", tls_error.c_str());
                ctx->warned_no_nvrtc = true;
            }
            use_jit = false;
            jit = nullptr;
        } else {
            blocks_per_sm = jit->kernel.blocks_per_sm;
        }
    }
    if (!use_jit) {
        shape = qe::TileShape();
        shape.steps = QE_INTERP_STEPS;
        shape.vec = QE_INTERP_VEC;
        blocks_per_sm = qe::interp_blocks_per_sm((int)mode);
    }
    const int64_t tile_rows = shape.rows_per_tile();
    const int block_threads = use_jit && (mode == qe::KM_FILTER || shape.pring) ? shape.block_threads() : QE_BLOCK;
    const int64_t ntiles64 = (n + tile_rows - 1) / tile_rows;
    const uint32_t ntiles = (uint32_t)ntiles64;
    if (ctx->opt_grid_per_sm) blocks_per_sm = std::min<int>(blocks_per_sm, (int)ctx->opt_grid_per_sm);
    if (mode == qe::KM_AGG) blocks_per_sm = std::min(blocks_per_sm, 8);
    // aggregation: CTAs take whole sub-tiles of QE_AGG_SUB_ROWS rows (qe_agg_body)
    if (mode == qe::KM_AGG && (QE_AGG_SUB_ROWS % tile_rows)) return fail(QE_ERR_STATE, "aggregation tile does not divide a sub-tile");
    const int64_t nsub = mode == qe::KM_AGG ? (ntiles64 + QE_AGG_SUB_ROWS / tile_rows - 1) / (QE_AGG_SUB_ROWS / tile_rows) : 0;
    const int64_t nsuper = (nsub + QE_AGG_SUPER_ROWS / QE_AGG_SUB_ROWS - 1) / (QE_AGG_SUPER_ROWS / QE_AGG_SUB_ROWS);
    int grid = (int)std::min<int64_t>(std::max<int64_t>(mode == qe::KM_AGG ? nsub : ntiles64, 1), (int64_t)ctx->num_sms * blocks_per_sm);

    // ---- result buffers (reused when the caller passes the previous result back)
    qe_result *r = *result;
    if (r && r->ctx != ctx) return fail(QE_ERR_INVALID_ARGUMENT, "result belongs to another context");
    const bool append = (d->flags & QE_RUN_APPEND) != 0;
    if (append) {
        if (mode == qe::KM_AGG) return fail(QE_ERR_UNSUPPORTED, "QE_RUN_APPEND: aggregations are combined with qe_agg_combine instead");
        if (!r || r->mode != (int)mode || r->cols.size() != bp.outs.size() || !r->ctl.ptr)
            return fail(QE_ERR_STATE, "QE_RUN_APPEND needs the result of an earlier run of the same pipeline");
    }
    const bool keep_pending = append || (extra && mode == qe::KM_AGG && extra->super_offset > 0);   // later chunk of one stream call
    if (r && r->pending && !keep_pending) { QE_CUDA(cudaEventSynchronize(r->ev_done)); r->pending = false; }
    if (r) fold_timing(r);
    if (!r) {
        r = new qe_result();
        r->ctx = ctx;
        cudaError_t e = cudaHostAlloc((void **)&r->host, sizeof(ResultHost), cudaHostAllocDefault);
        if (e == cudaSuccess) e = cudaEventCreate(&r->ev_begin);
        if (e == cudaSuccess) e = cudaEventCreate(&r->ev_end);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&r->ev_done, cudaEventDisableTiming);
        if (e != cudaSuccess) { qe_result_destroy(r); return fail_cuda("result setup", e); }
    }
    struct Cleanup {      // on error paths, do not leak a result we created
        qe_result *r; qe_result **slot; bool armed;
        ~Cleanup() { if (armed && *slot != r) qe_result_destroy(r); }
    } cleanup{r, result, true};

    const int64_t emitted_before = append ? r->batch_rows : 0;      // dense outputs: rows already in the result
    if (append && mode == qe::KM_PROJECT) {
        if (emitted_before % 128) return fail(QE_ERR_INVALID_ARGUMENT, "QE_RUN_APPEND: earlier batches must be multiples of 128 rows");
        if (emitted_before + ntiles64 * tile_rows > r->capacity)
            return fail(QE_ERR_CAPACITY, "QE_RUN_APPEND: the first run's out_capacity_rows does not cover this batch");
    }
    r->mode = (int)mode;
    if (mode == qe::KM_FILTER && r->f_sig != qkey) { r->f_sig = qkey; r->f_sel = -1.0; }
    r->used_jit = use_jit;
    r->batch_rows = emitted_before + n;
    r->agg_fns = bp.agg_fns;
    r->chained = extra && extra->chained;
    r->streamed = false;
    // capacity is fixed by the first run (out_capacity_rows may exceed the batch: later batches are appended)
    const int64_t capacity = append ? r->capacity
                           : mode == qe::KM_FILTER ? (d->out_capacity_rows > 0 ? d->out_capacity_rows : n)
                           : mode == qe::KM_PROJECT ? std::max(ntiles64 * tile_rows,
                                                               (d->out_capacity_rows + kPadRows - 1) / kPadRows * kPadRows + kPadRows)
                           : 0;
    r->capacity = capacity;
    if (r->cols.size() != bp.outs.size()) {
        for (ResultColumn &c : r->cols) { release(ctx, c.values); release(ctx, c.valid); }
        r->cols.assign(bp.outs.size(), ResultColumn());
    }

    QeParams p;
    memset(&p, 0, sizeof p);
    for (size_t s = 0; s < bp.slot_table_col.size(); s++) {
        const int c = bp.slot_table_col[s];
        const int type = table->cols[c].type;
        if (type == QE_TYPE_BOOLEAN) p.col_values[s] = (const char *)table->values[c] + row_begin / 8;
        else p.col_values[s] = (const char *)table->values[c] + (size_t)row_begin * type_width(type);
        p.col_valid[s] = table->valid[c] ? table->valid[c] + row_begin / 32 : nullptr;
    }
    p.nrows = n;
    p.ntiles = ntiles;
    p.sel_base = (uint32_t)(row_begin + d->sel_offset);
    p.out_capacity = capacity;
    p.out_base_mask = ~0ull;
    fill_literals(bp, p);

    const size_t bitmap_bytes = (size_t)((capacity + 31) / 32 + 16) * 4;
    if (mode != qe::KM_AGG) {
        for (size_t k = 0; k < bp.outs.size(); k++) {
            ResultColumn &c = r->cols[k];
            c.type = bp.outs[k].type;
            c.has_validity = bp.outs[k].has_validity;
            const size_t vbytes = c.type == QE_TYPE_BOOLEAN ? bitmap_bytes : (size_t)std::max<int64_t>(capacity, 1) * type_width(c.type);
            if ((rc = ensure(ctx, c.values, vbytes))) return rc;
            // dense outputs of an appended batch start where the earlier batches ended
            const size_t dense_off = mode == qe::KM_PROJECT ? (size_t)emitted_before : 0;
            p.out_values[k] = (char *)c.values.ptr + (c.type == QE_TYPE_BOOLEAN ? dense_off / 8 : dense_off * type_width(c.type));
            if (c.has_validity) {
                if ((rc = ensure(ctx, c.valid, bitmap_bytes))) return rc;
                p.out_valid[k] = (qe_u32 *)((char *)c.valid.ptr + dense_off / 8);
            }
            if (mode == qe::KM_FILTER && !append) {      // compacted bit columns are OR-ed into: start from zero
                if (c.type == QE_TYPE_BOOLEAN) QE_CUDA(cudaMemsetAsync(c.values.ptr, 0, bitmap_bytes, ctx->stream));
                if (c.has_validity) QE_CUDA(cudaMemsetAsync(c.valid.ptr, 0, bitmap_bytes, ctx->stream));
            }
        }
    } else {
        for (size_t k = 0; k < bp.outs.size(); k++) {
            r->cols[k].type = QE_TYPE_DOUBLE;       // every aggregate is DOUBLE, M/evaluator/TypeCheck.kt:110-120
            r->cols[k].has_validity = true;
        }
    }
    if (mode == qe::KM_FILTER && !(d->flags & QE_RUN_NO_SELECTION)) {
        if ((rc = ensure(ctx, r->sel, (size_t)std::max<int64_t>(capacity, 1) * 4))) return rc;
        p.sel = (qe_u32 *)r->sel.ptr;
    } else if (!append) {
        release(ctx, r->sel);                       // no selection vector this time: an earlier run's must not be readable
    }
    // control block: [0] kept-row count (u64)  [8] ticket (u32)  [16] append base (u64)  [32...] tile descriptors
    const size_t ctl_bytes = 32 + (mode == qe::KM_FILTER ? (size_t)ntiles * 8 : 0);
    if (append && r->ctl.bytes < ctl_bytes) {                 // grow, keeping the running count
        DevBuf bigger;
        if ((rc = ensure(ctx, bigger, ctl_bytes))) return rc;
        QE_CUDA(cudaMemcpyAsync(bigger.ptr, r->ctl.ptr, 32, cudaMemcpyDeviceToDevice, ctx->stream));
        release(ctx, r->ctl);                                  // stream-ordered: reused only by later work of this stream
        r->ctl = bigger;
    }
    if ((rc = ensure(ctx, r->ctl, ctl_bytes))) return rc;
    if (append) {
        // base <- count of the earlier batches; ticket and descriptors start from zero again
        QE_CUDA(cudaMemcpyAsync((char *)r->ctl.ptr + 16, r->ctl.ptr, 8, cudaMemcpyDeviceToDevice, ctx->stream));
        QE_CUDA(cudaMemsetAsync((char *)r->ctl.ptr + 8, 0, 8, ctx->stream));
        if (ctl_bytes > 32) QE_CUDA(cudaMemsetAsync((char *)r->ctl.ptr + 32, 0, ctl_bytes - 32, ctx->stream));
        p.out_base = (const qe_u64 *)((char *)r->ctl.ptr + 16);
    } else {
        QE_CUDA(cudaMemsetAsync(r->ctl.ptr, 0, ctl_bytes, ctx->stream));
        if (extra && extra->chained && mode == qe::KM_FILTER) {
            // a streamed chunk with buffers of its own: positions start at (rows kept so far) & 7, so that compacted bit
            // columns keep their phase and the chunks' bytes can be laid end to end on the host; the kept-row count the
            // kernel leaves in ctl[0] is still the running total
            if (extra->chain_base) {
                QE_CUDA(cudaMemcpyAsync((char *)r->ctl.ptr + 16, extra->chain_base, 8, cudaMemcpyDeviceToDevice, ctx->stream));
                QE_CUDA(cudaMemcpyAsync(r->ctl.ptr, extra->chain_base, 8, cudaMemcpyDeviceToDevice, ctx->stream));   // n == 0: the total stands
                p.out_base = (const qe_u64 *)((char *)r->ctl.ptr + 16);
            }
            p.out_base_mask = 7ull;
        }
    }
    p.out_count = (qe_u64 *)r->ctl.ptr;
    p.ticket = (qe_u32 *)((char *)r->ctl.ptr + 8);
    p.tile_state = (qe_u64 *)((char *)r->ctl.ptr + 32);
    const int nagg = (int)bp.agg_fns.size();
    const int64_t super_off = extra ? extra->super_offset : 0;
    if (mode == qe::KM_AGG) {
        if ((rc = ensure(ctx, r->partials, (size_t)std::max<int64_t>(nsub, 1) * (size_t)nagg * 4 * sizeof(double)))) return rc;
        if ((rc = ensure(ctx, r->final_aggs, (size_t)std::max<int64_t>(nsuper, 1) * (size_t)nagg * 4 * sizeof(double)))) return rc;
        const int64_t total = extra && extra->super_total > 0 ? extra->super_total : nsuper;
        if (super_off + nsuper > total) return fail(QE_ERR_STATE, "super-tile partial list too short");
        if ((rc = ensure_host_super(r, (size_t)std::max<int64_t>(total, 1) * (size_t)nagg * 4))) return rc;
        r->n_super = super_off + nsuper;
        r->agg_folded = false;
        p.agg_partials = (double *)r->partials.ptr;
    }
    if (ctx->opt_debug_flags & 16) {
        if ((rc = ensure(ctx, r->debug, 4096 * 8 * sizeof(uint64_t)))) return rc;
        QE_CUDA(cudaMemsetAsync(r->debug.ptr, 0, 4096 * 8 * sizeof(uint64_t), ctx->stream));
        p.debug = (qe_u64 *)r->debug.ptr;
    }
    DevBuf iprog;
    if (!use_jit) {
        std::vector<uint32_t> words;
        serialize_interp(bp, words);
        if ((rc = ensure(ctx, iprog, words.size() * 4))) return rc;
        // pageable source: the runtime stages it before returning, so `words` may go out of scope
        QE_CUDA(cudaMemcpyAsync(iprog.ptr, words.data(), words.size() * 4, cudaMemcpyHostToDevice, ctx->stream));
        p.iprog = (const qe_u32 *)iprog.ptr;
    }

    // ---- launch
    QE_CUDA(cudaEventRecord(r->ev_begin, ctx->stream));
    if (n > 0) {
        if (use_jit) {
            void *args[] = {&p};
            QE_CUDA(cudaLaunchKernel((const void *)jit->kernel.kernel, dim3(grid), dim3(block_threads), args, jit->kernel.dyn_smem, ctx->stream));
        } else {
            QE_CUDA(qe::launch_interp((int)mode, p, grid, ctx->stream));
        }
        ctx->launches++;
    }
    QE_CUDA(cudaEventRecord(r->ev_end, ctx->stream));
    if (mode == qe::KM_AGG && n > 0) {
        // second level of the fixed-shape fold on the device, the (few) super-tile partials to the host list
        QE_CUDA(qe::launch_agg_final((const double *)r->partials.ptr, nsub, nagg, (double *)r->final_aggs.ptr, ctx->stream));
        ctx->launches++;
        QE_CUDA(cudaMemcpyAsync(r->h_super + (size_t)super_off * (size_t)nagg * 4, r->final_aggs.ptr,
                                (size_t)nsuper * (size_t)nagg * 4 * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    }
    if (extra && extra->chained)     // inside a stream the copy engine is busy with the downloads: see launch_store_u64
        QE_CUDA(qe::launch_store_u64(&r->host->count, (const unsigned long long *)r->ctl.ptr, ctx->stream));
    else
        QE_CUDA(cudaMemcpyAsync(&r->host->count, r->ctl.ptr, sizeof(unsigned long long), cudaMemcpyDeviceToHost, ctx->stream));
    QE_CUDA(cudaEventRecord(r->ev_done, ctx->stream));
    r->pending = true;
    r->timed = true;
    if (iprog.ptr) {
        // returned to the pool only after the stream has consumed it
        cudaStreamSynchronize(ctx->stream);
        release(ctx, iprog);
    }
    cleanup.armed = false;
    *result = r;
    return QE_OK;
}

extern "C" int qe_pipeline_run_ex(qe_ctx *ctx, const qe_table *table, const qe_pipeline_desc *d, qe_result **result)
{
    return run_pipeline(ctx, table, d, result, nullptr);
}

extern "C" int qe_pipeline_run(qe_ctx *ctx, const qe_table *table, const qe_program *filter,
                               const qe_program *const *proj, int nproj, const int32_t *agg_fns, int nagg,
                               qe_result **result)
{
    qe_pipeline_desc d;
    memset(&d, 0, sizeof d);
    d.filter = filter;
    d.proj = proj;
    d.nproj = nproj;
    d.agg_fns = agg_fns;
    d.nagg = nagg;
    d.row_begin = 0;
    d.row_end = -1;
    return qe_pipeline_run_ex(ctx, table, &d, result);
}


// ------------------------------------------------------------------ group-by
// Stands in for Scan -> [Filter] -> Projection(keys + aggregate inputs) -> GroupByAggregationOperator
// (M/evaluator/Planner.kt:46-58, M/operator/GroupByAggregationOperator.kt:21-75).  Synchronous: the groups
// are few compared with the rows, they are finished (Accumulator.finish) and ordered by first appearance
// (LinkedHashMap iteration order) on the host.

static void finish_agg(int fn, const qe_agg_partial &p, double *value, int *is_null);

static int groupby_finish(qe_ctx *ctx, qe_result *r, const qe::BoundPipeline &bp, const qe::GroupPlan &gp, int64_t ngroups)
{
    const int nkeys = gp.nkeys, nagg = (int)bp.agg_fns.size(), nwords = gp.nwords;
    const size_t rec = (size_t)nwords + (size_t)nkeys + 1;
    std::vector<uint64_t> records((size_t)ngroups * rec);
    if (ngroups > 0) {
        QE_CUDA(cudaMemcpyAsync(records.data(), r->g_records.ptr, records.size() * 8, cudaMemcpyDeviceToHost, ctx->stream));
        QE_CUDA(cudaStreamSynchronize(ctx->stream));
    }
    std::vector<int64_t> order((size_t)ngroups);
    for (int64_t i = 0; i < ngroups; i++) order[(size_t)i] = i;
    // word 0 = first row << 32 | rows: ordering by the whole word orders by first row (first rows are distinct)
    std::sort(order.begin(), order.end(), [&](int64_t a, int64_t b) { return records[(size_t)a * rec] < records[(size_t)b * rec]; });

    const int ncols = nkeys + nagg;
    r->h_rows = ngroups;
    r->h_values.assign((size_t)ncols, std::vector<uint8_t>());
    r->h_valid.assign((size_t)ncols, std::vector<uint8_t>());
    r->g_partials.assign((size_t)ngroups * (size_t)std::max(nagg, 1), qe_agg_partial());
    r->g_first_row.assign((size_t)ngroups, 0);
    const size_t vbytes = (size_t)((ngroups + 7) / 8);
    for (int c = 0; c < ncols; c++) {
        const int type = r->cols[(size_t)c].type;
        r->h_values[(size_t)c].assign(type == QE_TYPE_BOOLEAN ? vbytes : (size_t)ngroups * type_width(type), 0);
        r->h_valid[(size_t)c].assign(vbytes, 0);
    }
    for (int64_t g = 0; g < ngroups; g++) {
        const uint64_t *w = &records[(size_t)order[(size_t)g] * rec];
        r->g_first_row[(size_t)g] = (int64_t)(w[0] >> 32);
        const uint32_t nulls = (uint32_t)w[nwords + nkeys];
        for (int k = 0; k < nkeys; k++) {
            if ((nulls >> k) & 1u) continue;                    // null key: validity bit stays 0, value 0
            r->h_valid[(size_t)k][(size_t)(g / 8)] |= (uint8_t)(1u << (g % 8));
            const uint64_t bits = w[nwords + k];
            uint8_t *dst = r->h_values[(size_t)k].data();
            switch (r->cols[(size_t)k].type) {
            case QE_TYPE_BOOLEAN: if (bits) dst[g / 8] |= (uint8_t)(1u << (g % 8)); break;
            case QE_TYPE_DOUBLE: case QE_TYPE_INT64: memcpy(dst + (size_t)g * 8, &bits, 8); break;
            default: { const int32_t v = (int32_t)(int64_t)bits; memcpy(dst + (size_t)g * 4, &v, 4); break; }
            }
        }
        for (int j = 0; j < nagg; j++) {
            qe_agg_partial p;
            p.sum = 0.0;
            p.min = 0.0;
            p.max = 0.0;
            p.count = gp.cnt_word[(size_t)j] < 0 ? (int64_t)(w[0] & 0xffffffffULL) : (int64_t)w[gp.cnt_word[(size_t)j]];
            const int fn = bp.agg_fns[(size_t)j];
            if (gp.val_word[(size_t)j] >= 0) {
                double v;
                memcpy(&v, &w[gp.val_word[(size_t)j]], 8);
                if (fn == QE_AGG_FN_MIN) p.min = v;
                else if (fn == QE_AGG_FN_MAX) p.max = v;
                else p.sum = v;
            }
            r->g_partials[(size_t)g * (size_t)nagg + (size_t)j] = p;
            double value = 0.0;
            int is_null = 0;
            finish_agg(fn, p, &value, &is_null);
            if (!is_null) {
                memcpy(r->h_values[(size_t)(nkeys + j)].data() + (size_t)g * 8, &value, 8);
                r->h_valid[(size_t)(nkeys + j)][(size_t)(g / 8)] |= (uint8_t)(1u << (g % 8));
            }
        }
    }
    return QE_OK;
}


extern "C" int qe_groupby_run(qe_ctx *ctx, const qe_table *table, const qe_groupby_desc *d, qe_result **result)
{
    if (!ctx || !table || !d || !result) return fail(QE_ERR_INVALID_ARGUMENT, "NULL argument");
    if (table->ctx != ctx) return fail(QE_ERR_INVALID_ARGUMENT, "table belongs to another context");
    if (d->nkeys < 0 || d->nagg < 0 || d->nkeys + d->nagg <= 0) return fail(QE_ERR_INVALID_ARGUMENT, "bad key / aggregate count");
    if ((d->nkeys && !d->keys) || (d->nagg && (!d->inputs || !d->agg_fns))) return fail(QE_ERR_INVALID_ARGUMENT, "NULL key / input list");
    DeviceGuard g(ctx->device);

    std::vector<qe::ColumnInfo> schema(table->cols.size());
    for (size_t c = 0; c < table->cols.size(); c++) {
        schema[c].type = table->cols[c].type;
        schema[c].nullable = table->cols[c].nullable != 0;
    }
    std::vector<const qe::Program *> proj;
    for (int k = 0; k < d->nkeys; k++) {
        if (!d->keys[k]) return fail(QE_ERR_INVALID_ARGUMENT, "NULL key program");
        proj.push_back(&d->keys[k]->prog);
    }
    for (int k = 0; k < d->nagg; k++) {
        if (!d->inputs[k]) return fail(QE_ERR_INVALID_ARGUMENT, "NULL aggregate input program");
        proj.push_back(&d->inputs[k]->prog);
    }
    std::vector<int> aggs(d->agg_fns, d->agg_fns + d->nagg);
    qe::BoundPipeline bp;
    std::string err;
    int rc = qe::bind_pipeline(schema, d->filter ? &d->filter->prog : nullptr, proj, aggs, bp, err, d->nkeys);
    if (rc) return fail(rc, err);
    // Variants, tried in this order until one holds every group a CTA meets: private accumulators with 512 threads
    // (16 warps per SM hide the shared-memory and HBM latencies; half as many groups fit), with 256 threads, atomics.
    const qe::GroupPlan gp512 = qe::plan_groupby(bp, 512, (int)ctx->opt_lane_acc_kb);
    const qe::GroupPlan gp = qe::plan_groupby(bp, 256);
    if (gp.nwords > 48) return fail(QE_ERR_UNSUPPORTED, "too many accumulator words");

    const int64_t row_begin = d->row_begin;
    const int64_t row_end = d->row_end < 0 ? table->nrows : d->row_end;
    if (row_begin < 0 || row_end < row_begin || row_end > table->nrows) return fail(QE_ERR_INVALID_ARGUMENT, "bad row range");
    if (row_begin % 128) return fail(QE_ERR_INVALID_ARGUMENT, "row_begin must be a multiple of 128");
    const int64_t n = row_end - row_begin;

    const int64_t max_groups = d->max_groups > 0 ? d->max_groups : 65536;
    if (max_groups > (1ll << 28)) return fail(QE_ERR_INVALID_ARGUMENT, "max_groups too large");
    int64_t nslots = 1024;
    while (nslots < 2 * max_groups) nslots *= 2;

    qe_result *r = *result;
    if (r && r->ctx != ctx) return fail(QE_ERR_INVALID_ARGUMENT, "result belongs to another context");
    if (r && r->pending) { QE_CUDA(cudaEventSynchronize(r->ev_done)); r->pending = false; }
    if (r) fold_timing(r);
    if (!r) {
        r = new qe_result();
        r->ctx = ctx;
        cudaError_t e = cudaHostAlloc((void **)&r->host, sizeof(ResultHost), cudaHostAllocDefault);
        if (e == cudaSuccess) e = cudaEventCreate(&r->ev_begin);
        if (e == cudaSuccess) e = cudaEventCreate(&r->ev_end);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&r->ev_done, cudaEventDisableTiming);
        if (e != cudaSuccess) { qe_result_destroy(r); return fail_cuda("result setup", e); }
    }
    struct Cleanup {
        qe_result *r; qe_result **slot; bool armed;
        ~Cleanup() { if (armed && *slot != r) qe_result_destroy(r); }
    } cleanup{r, result, true};

    for (ResultColumn &c : r->cols) { release(ctx, c.values); release(ctx, c.valid); }
    r->cols.assign(bp.outs.size(), ResultColumn());
    for (size_t k = 0; k < bp.outs.size(); k++) {
        r->cols[k].type = k < (size_t)gp.nkeys ? bp.outs[k].type : QE_TYPE_DOUBLE;
        r->cols[k].has_validity = true;
    }
    r->agg_fns = bp.agg_fns;
    r->g_nkeys = gp.nkeys;
    r->batch_rows = n;
    r->capacity = 0;
    r->used_jit = true;
    r->h_rows = 0;

    qe::GroupWords gw;
    memset(&gw, 0, sizeof gw);
    gw.nwords = gp.nwords;
    for (int w = 0; w < gp.nwords; w++) { gw.kind[w] = (unsigned char)gp.kind[(size_t)w]; gw.init[w] = gp.init[(size_t)w]; }

    const int nk = std::max(gp.nkeys, 1);
    if ((rc = ensure(ctx, r->g_state, (size_t)nslots * 4))) return rc;
    if ((rc = ensure(ctx, r->g_keys, (size_t)nslots * (size_t)nk * 8))) return rc;
    if ((rc = ensure(ctx, r->g_nulls, (size_t)nslots * 4))) return rc;
    if ((rc = ensure(ctx, r->g_acc, (size_t)nslots * (size_t)gp.nwords * 8))) return rc;
    if ((rc = ensure(ctx, r->g_ctl, 16))) return rc;

    QeParams p;
    memset(&p, 0, sizeof p);
    for (size_t s = 0; s < bp.slot_table_col.size(); s++) {
        const int c = bp.slot_table_col[s];
        const int type = table->cols[c].type;
        if (type == QE_TYPE_BOOLEAN) p.col_values[s] = (const char *)table->values[c] + row_begin / 8;
        else p.col_values[s] = (const char *)table->values[c] + (size_t)row_begin * type_width(type);
        p.col_valid[s] = table->valid[c] ? table->valid[c] + row_begin / 32 : nullptr;
    }
    p.nrows = n;
    p.sel_base = (uint32_t)row_begin;
    p.g_state = (qe_u32 *)r->g_state.ptr;
    p.g_keys = (qe_u64 *)r->g_keys.ptr;
    p.g_nulls = (qe_u32 *)r->g_nulls.ptr;
    p.g_acc = (qe_u64 *)r->g_acc.ptr;
    p.g_ctl = (qe_u32 *)r->g_ctl.ptr;
    p.g_slot_mask = (qe_u32)(nslots - 1);
    p.g_max_groups = (qe_u32)max_groups;
    p.out_base_mask = ~0ull;
    fill_literals(bp, p);

    uint32_t ctl[4] = {0, 0, 0, 0};
    // a result that is reused remembers which variants this group-by has overflowed
    const std::string sig = query_key(bp);
    // levels: 0 registers, 1 private 512 threads, 2 private 256 threads, 3 lane copies (32 per CTA), 4 eight copies,
    // 5 one copy per warp (256 threads, no atomics), 6 two copies / one copy, 7 two passes (records partitioned by key
    // hash, one CTA per partition), 8 global atomics
    int level = r->g_overflowed == sig ? r->g_level : 0;
    if (d->flags & QE_GROUPBY_ATOMIC) level = 8;
    else if (d->flags & QE_GROUPBY_TWOPASS) level = std::max(level, 7);
    else if (d->flags & QE_GROUPBY_SHARED) level = std::max(level, 3);
    // register variant: every group's words are touched for every row, so compile for the group count the same
    // group-by found last time (a reused result remembers it); more groups than that -> once more with the maximum
    const int rmax_full = ctx->opt_groupby_registers == 2 ? gp.rmax_regs : gp.rmax;
    int rmax_use = rmax_full;
    if (r->g_seen_sig == sig && r->g_seen_groups > 0) rmax_use = std::min(rmax_full, std::max(4, r->g_seen_groups));
    // the most copies per CTA that hold the groups the same group-by found last time (every CTA meets nearly all of them)
    int hint_copies = 0;
    if (r->g_seen_sig == sig && r->g_seen_groups > 0)
        for (int c = 32; c >= 1 && !hint_copies; c /= 2)
            if (gp512.clmax[qe::GroupPlan::copies_index(c)] >= r->g_seen_groups) hint_copies = c;
    for (;; level++) {
        const bool regs = level == 0;
        qe::TileShape reg_shape;
        if (regs) {
            if (!ctx->opt_groupby_registers || rmax_full < 2) continue;
            reg_shape = choose_shape(ctx, bp, qe::KM_GROUPBY_REG);
            reg_shape.rmax_hint = rmax_use;
            // the final reduce lays the thread copies out over the stages
            if (!reg_shape.greg || (size_t)gp.nwords * (size_t)rmax_use * (size_t)reg_shape.ncw * 32u * 8u > qe::staged_smem_bytes(bp, reg_shape)) continue;
        }
        if (level == 1 && gp512.lmax < 2) continue;
        if (level == 2 && (gp.lmax < 1 || gp.lmax <= gp512.lmax)) continue;
        // few words per group: the lane copies are faster than the 256-thread private copies (SUM + COUNT over 16-byte
        // rows: 2.2 ms against 2.5 ms at 32-48 groups) and hold more groups
        if (level == 2 && gp.nwords <= 3 && ctx->opt_groupby_lanes && gp512.clmax[5] >= gp.lmax) continue;
        // shared-memory copies: level 3 = 32 per CTA, 4 = 8, 6 = 2 (1 where one copy holds more groups); a group-by that knows how many groups it found last
        // time goes straight to the most copies that hold them (16 and 4 included)
        int copies = level == 3 ? 32 : level == 4 ? 8 : level == 6 ? (gp512.clmax[0] > gp512.clmax[1] ? 1 : 2) : 0;
        if (copies) {
            if (!ctx->opt_groupby_lanes && copies > 2) continue;
            if (hint_copies > 0 && copies > hint_copies) continue;                       // would overflow again
            if (hint_copies > copies) copies = hint_copies;
            if (gp512.clmax[qe::GroupPlan::copies_index(copies)] < 16) continue;
        }
        if (level == 5 && (gp.slmaxw < 64 || !ctx->opt_groupby_warps)) continue;
        if (level == 7) {
            if (!ctx->opt_groupby_twopass || gp512.clmax[0] < 64 || (n < (1 << 22) && !(d->flags & QE_GROUPBY_TWOPASS))) continue;      // small batches: atomics do
            // pass 2 holds a partition's share of the groups: the most copies that do, by the count of the last run
            // (the groups spread evenly over the partitions: + 25 % and a few dozen), else one copy (the most groups)
            int np = 0, pc = 0;
            qe::GroupPlan::scatter_shape(gp512.rec_words, np, pc);
            int copies2 = gp512.clmax[0] >= gp512.clmax[1] ? 1 : 2;
            if (r->g_seen_sig == sig && r->g_seen_groups > 0) {
                const int64_t per_part = (int64_t)r->g_seen_groups / np * 5 / 4 + 48;
                for (int c = 32; c > copies2; c /= 2)
                    if (gp512.clmax[qe::GroupPlan::copies_index(c)] >= per_part) { copies2 = c; break; }
            }
            qe::TileShape s1 = choose_shape(ctx, bp, qe::KM_GROUPBY_SHARED);
            s1.gthreads = 512;
            s1.steps = 1;
            s1.vec = qe::jit_wide_loads_supported() ? 4 : 2;
            s1.lane_acc_kb = (int)ctx->opt_lane_acc_kb;
            s1.npart = np;
            s1.pcap = pc;
            s1.scopies = copies2;
            qe::TileShape s2 = s1;
            s1.gpass = 1;
            s2.gpass = 2;
            JitEntry *j1 = nullptr, *j2 = nullptr;
            rc = get_jit_kernel(ctx, bp, qe::KM_GROUPBY_SHARED, s1, &j1, qe::GroupPlan::scatter_smem(gp512.rec_words, np, pc));
            if (rc < 0) return fail(QE_ERR_JIT, "group-by needs generated code: " + tls_error);
            if (rc) return rc;
            rc = get_jit_kernel(ctx, bp, qe::KM_GROUPBY_SHARED, s2, &j2, gp512.copies_smem(copies2));
            if (rc < 0) return fail(QE_ERR_JIT, "group-by needs generated code: " + tls_error);
            if (rc) return rc;
            const int64_t tile_rows = s1.rows_per_tile();
            const int64_t ntiles64 = (n + tile_rows - 1) / tile_rows;
            p.ntiles = (uint32_t)ntiles64;
            // Space per partition: the rows spread like the groups do -- with g groups per partition the largest of the
            // partitions holds about (1 + 4.5 / sqrt(g)) times the mean.  g from the count of the last run, else from
            // the groups the table held when the last variant gave up (a lower bound).
            const double g_est = (double)std::max<int64_t>(r->g_seen_sig == sig ? r->g_seen_groups : 0, std::max<int64_t>(ctl[0], gp512.clmax[0])) / np;
            const double factor = std::min(3.0, 1.08 + 4.5 / std::sqrt(std::max(g_est, 1.0)));
            // (a small batch -- only QE_GROUPBY_TWOPASS sends one here -- gets room for all its rows in any one partition)
            const int64_t pt_cap = (std::max<int64_t>((int64_t)((double)(n / np) * factor) + 4096, std::min<int64_t>(n, 1 << 16)) + 31) / 32 * 32;
            if (pt_cap > 0xffffffffll) continue;
            if ((rc = ensure(ctx, r->g_pt_words, (size_t)gp512.rec_words * (size_t)np * (size_t)pt_cap * 8))) return rc;
            if ((rc = ensure(ctx, r->g_pt_cursor, (size_t)np * 4))) return rc;
            p.pt_words = (qe_u64 *)r->g_pt_words.ptr;
            p.pt_cursor = (qe_u32 *)r->g_pt_cursor.ptr;
            p.pt_cap = (qe_u32)pt_cap;
            r->mode = (int)qe::KM_GROUPBY_SHARED;
            r->g_variant = 7;
            r->g_copies = copies2;
            r->g_threads = 512;
            QE_CUDA(qe::launch_groupby_init((uint32_t *)r->g_state.ptr, (uint64_t *)r->g_acc.ptr, (uint32_t *)r->g_ctl.ptr, nslots, gw, ctx->stream));
            QE_CUDA(cudaMemsetAsync(r->g_pt_cursor.ptr, 0, (size_t)np * 4, ctx->stream));
            ctx->launches++;
            QE_CUDA(cudaEventRecord(r->ev_begin, ctx->stream));
            void *args[] = {&p};
            const int grid1 = (int)std::min<int64_t>(std::max<int64_t>(ntiles64, 1), (int64_t)ctx->num_sms);
            QE_CUDA(cudaLaunchKernel((const void *)j1->kernel.kernel, dim3(grid1), dim3(512), args, j1->kernel.dyn_smem, ctx->stream));
            QE_CUDA(cudaLaunchKernel((const void *)j2->kernel.kernel, dim3(std::min(np, ctx->num_sms)), dim3(512), args, j2->kernel.dyn_smem, ctx->stream));
            ctx->launches += 2;
            QE_CUDA(cudaEventRecord(r->ev_end, ctx->stream));
            r->timed = true;
            QE_CUDA(cudaMemcpyAsync(ctl, r->g_ctl.ptr, sizeof ctl, cudaMemcpyDeviceToHost, ctx->stream));
            QE_CUDA(cudaStreamSynchronize(ctx->stream));
            if (ctl[1] & 2u)
                return fail(QE_ERR_CAPACITY, "more than max_groups (" + std::to_string(max_groups) + ") groups; rerun with a larger max_groups");
            if (!(ctl[1] & 1u)) break;
            fold_timing(r);      // a partition outgrew its space or its CTA's table (one key with a large share of the rows): atomics
            r->g_overflowed = sig;
            r->g_level = level + 1;
            continue;
        }
        const bool atomic = level >= 8, shared = level >= 3 && level <= 6, warpc = level == 5;
        const qe::GroupPlan &plan = (level == 1 || (shared && !warpc)) ? gp512 : gp;
        const qe::KernelMode mode = regs ? qe::KM_GROUPBY_REG : atomic ? qe::KM_GROUPBY_ATOMIC : shared ? qe::KM_GROUPBY_SHARED : qe::KM_GROUPBY_PRIVATE;
        const int ci = qe::GroupPlan::copies_index(copies ? copies : 1);
        const int part_groups = regs ? rmax_use : warpc ? plan.slmaxw : shared ? plan.clmax[ci] : plan.lmax;
        qe::TileShape shape = regs ? reg_shape : choose_shape(ctx, bp, mode);
        shape.scopies = warpc ? -1 : copies ? copies : 1;
        shape.lane_acc_kb = (int)ctx->opt_lane_acc_kb;
        if (!atomic && !regs) {
            shape.gthreads = plan.threads;
            if (plan.threads == 512 && !ctx->opt_steps && !ctx->opt_vec) {      // measured best for 16- and 48-byte rows
                shape.steps = 1;
                shape.vec = qe::jit_wide_loads_supported() ? 4 : 2;
            }
        }
        // a tile (and the prefetch of the next one) must stay inside the padding of the column allocations
        if (shape.rows_per_tile() > kPadRows)
            return fail(QE_ERR_INVALID_ARGUMENT, "group-by tile of " + std::to_string(shape.rows_per_tile()) + " rows (threads / 32 * steps * 32 * vec) exceeds " +
                                                 std::to_string(kPadRows) + ": lower the \"steps\" / \"vec\" options");
        JitEntry *jit = nullptr;
        rc = get_jit_kernel(ctx, bp, mode, shape, &jit, (atomic || regs) ? 0 : warpc ? plan.shared_smemw() : shared ? plan.copies_smem(copies) : plan.private_smem());
        if (rc < 0) return fail(QE_ERR_JIT, "group-by needs generated code: " + tls_error);
        if (rc) return rc;
        const int64_t tile_rows = shape.rows_per_tile();
        const int64_t ntiles64 = (n + tile_rows - 1) / tile_rows;
        int blocks_per_sm = jit->kernel.blocks_per_sm;
        if (ctx->opt_grid_per_sm) blocks_per_sm = std::min<int>(blocks_per_sm, (int)ctx->opt_grid_per_sm);
        const int grid = (int)std::min<int64_t>(std::max<int64_t>(ntiles64, 1), (int64_t)ctx->num_sms * blocks_per_sm);
        p.ntiles = (uint32_t)ntiles64;
        if (!atomic) {
            if ((rc = ensure(ctx, r->g_part, (size_t)grid * (size_t)part_groups * (size_t)plan.nwords * 8))) return rc;
            if ((rc = ensure(ctx, r->g_part_slot, (size_t)grid * (size_t)(1 + part_groups) * 4))) return rc;
            p.g_part = (qe_u64 *)r->g_part.ptr;
            p.g_part_slot = (qe_u32 *)r->g_part_slot.ptr;
        }
        r->mode = (int)mode;
        r->g_variant = regs ? 3 : atomic ? 1 : warpc ? 6 : copies == 32 ? 4 : copies >= 8 ? 5 : shared ? 2 : 0;
        r->g_copies = shared && !warpc ? copies : 0;
        const int threads = regs ? shape.block_threads() : atomic ? QE_BLOCK : plan.threads;
        r->g_threads = threads;
        QE_CUDA(qe::launch_groupby_init((uint32_t *)r->g_state.ptr, (uint64_t *)r->g_acc.ptr, (uint32_t *)r->g_ctl.ptr, nslots, gw, ctx->stream));
        ctx->launches++;
        QE_CUDA(cudaEventRecord(r->ev_begin, ctx->stream));
        if (n > 0) {
            void *args[] = {&p};
            QE_CUDA(cudaLaunchKernel((const void *)jit->kernel.kernel, dim3(grid), dim3(threads), args,
                                     jit->kernel.dyn_smem, ctx->stream));
            ctx->launches++;
        }
        QE_CUDA(cudaEventRecord(r->ev_end, ctx->stream));
        r->timed = true;
        if (n > 0 && !atomic) {
            QE_CUDA(qe::launch_groupby_merge((const uint64_t *)r->g_part.ptr, (const uint32_t *)r->g_part_slot.ptr, (uint64_t *)r->g_acc.ptr,
                                             grid, part_groups, gw, ctx->stream));
            ctx->launches++;
        }
        QE_CUDA(cudaMemcpyAsync(ctl, r->g_ctl.ptr, sizeof ctl, cudaMemcpyDeviceToHost, ctx->stream));
        QE_CUDA(cudaStreamSynchronize(ctx->stream));
        if (ctl[1] & 2u)
            return fail(QE_ERR_CAPACITY, "more than max_groups (" + std::to_string(max_groups) + ") groups; rerun with a larger max_groups");
        if (!(ctl[1] & 1u)) break;
        if (atomic) return fail(QE_ERR_STATE, "group-by: unexpected overflow flag");
        fold_timing(r);          // the discarded attempt still ran: keep its time in the totals
        if (regs && rmax_use < rmax_full) {                    // more groups than last time: the same variant at full size
            rmax_use = rmax_full;
            level--;
            continue;
        }
        r->g_overflowed = sig;   // a CTA met more groups than fit its shared memory: next variant
        r->g_level = level + 1;
        hint_copies = 0;         // the hint was wrong (other data): the plain ladder from here
    }
    const int64_t ngroups = ctl[0];
    r->g_seen_sig = sig;
    r->g_seen_groups = (int)std::min<int64_t>(ngroups, 1 << 20);
    if (ngroups > 0) {
        if ((rc = ensure(ctx, r->g_records, (size_t)ngroups * (size_t)(gp.nwords + gp.nkeys + 1) * 8))) return rc;
        QE_CUDA(qe::launch_groupby_collect((const uint32_t *)r->g_state.ptr, (const uint64_t *)r->g_keys.ptr, (const uint32_t *)r->g_nulls.ptr,
                                           (const uint64_t *)r->g_acc.ptr, nslots, gp.nkeys, gp.nwords, (uint64_t *)r->g_records.ptr,
                                           (uint32_t *)r->g_ctl.ptr, (uint32_t)ngroups, ctx->stream));
        ctx->launches++;
    }
    if ((rc = groupby_finish(ctx, r, bp, gp, ngroups))) return rc;
    QE_CUDA(cudaEventRecord(r->ev_done, ctx->stream));
    r->pending = false;
    r->host->count = (unsigned long long)ngroups;
    cleanup.armed = false;
    *result = r;
    return QE_OK;
}

extern "C" int qe_result_group_partial(qe_result *r, int64_t group, int k, qe_agg_partial *out)
{
    if (!r || !out) return fail(QE_ERR_INVALID_ARGUMENT, "NULL argument");
    if (!is_groupby(r)) return fail(QE_ERR_STATE, "this result is not a group-by");
    const int nagg = (int)r->agg_fns.size();
    if (group < 0 || group >= r->h_rows || k < 0 || k >= nagg) return fail(QE_ERR_INVALID_ARGUMENT, "group / aggregate index out of range");
    *out = r->g_partials[(size_t)group * (size_t)nagg + (size_t)k];
    return QE_OK;
}

extern "C" int qe_result_group_first_rows(qe_result *r, int64_t *dst)
{
    if (!r || !dst) return fail(QE_ERR_INVALID_ARGUMENT, "NULL argument");
    if (!is_groupby(r)) return fail(QE_ERR_STATE, "this result is not a group-by");
    for (int64_t g = 0; g < r->h_rows; g++) dst[g] = r->g_first_row[(size_t)g];
    return QE_OK;
}

extern "C" int qe_result_group_variant(const qe_result *r)
{
    return r && is_groupby(r) ? r->g_variant : -1;
}

extern "C" int qe_result_group_copies(const qe_result *r)
{
    return r && is_groupby(r) ? r->g_copies : -1;
}

extern "C" int qe_result_group_threads(const qe_result *r)
{
    return r && is_groupby(r) ? r->g_threads : -1;
}

// ------------------------------------------------------- result accessors

static int wait_result(qe_result *r)
{
    if (!r) return fail(QE_ERR_INVALID_ARGUMENT, "result is NULL");
    if (r->pending) {
        DeviceGuard g(r->ctx->device);
        QE_CUDA(cudaEventSynchronize(r->ev_done));
        r->pending = false;
    }
    if (r->mode == qe::KM_FILTER && !r->chained && !r->streamed && r->batch_rows > 0 && !r->f_sig.empty() &&
        (int64_t)r->host->count <= r->capacity) {
        if (r->ctx->sel_memo.size() > 4096) r->ctx->sel_memo.clear();
        r->ctx->sel_memo[r->f_sig] = (double)r->host->count / (double)r->batch_rows;
    }
    fold_aggs(r);
    return QE_OK;
}

static int64_t emitted_rows(const qe_result *r)
{
    if (is_groupby(r)) return r->h_rows;
    if (r->mode == qe::KM_AGG) return 1;
    if (r->mode == qe::KM_PROJECT) return r->batch_rows;
    return (int64_t)r->host->count;
}

extern "C" int qe_result_num_rows(qe_result *r, int64_t *nrows)
{
    int rc = wait_result(r);
    if (rc) return rc;
    if (nrows) *nrows = emitted_rows(r);
    if (r->mode == qe::KM_FILTER && !r->chained && !r->streamed && (int64_t)r->host->count > r->capacity) {
        return fail(QE_ERR_CAPACITY, "filter kept " + std::to_string(r->host->count) + " rows but out_capacity_rows is " +
                                     std::to_string(r->capacity) + "; rerun with a larger capacity");
    }
    return QE_OK;
}

extern "C" int qe_result_num_columns(const qe_result *r, int *ncols)
{
    if (!r || !ncols) return fail(QE_ERR_INVALID_ARGUMENT, "NULL argument");
    *ncols = (int)r->cols.size();
    return QE_OK;
}

extern "C" int qe_result_column_type(const qe_result *r, int k, int *type, int *has_validity)
{
    if (!r || k < 0 || k >= (int)r->cols.size()) return fail(QE_ERR_INVALID_ARGUMENT, "column index out of range");
    if (type) *type = r->cols[k].type;
    if (has_validity) *has_validity = r->cols[k].has_validity ? 1 : 0;
    return QE_OK;
}

static void fill_valid_bitmap(uint8_t *dst, int64_t n)
{
    memset(dst, 0xff, (size_t)(n / 8));
    if (n % 8) dst[n / 8] = (uint8_t)((1u << (n % 8)) - 1u);
}

extern "C" int qe_result_column(qe_result *r, int k, void *dst_values, uint8_t *dst_validity)
{
    int64_t n = 0;
    int rc = qe_result_num_rows(r, &n);
    if (rc) return rc;
    if (k < 0 || k >= (int)r->cols.size()) return fail(QE_ERR_INVALID_ARGUMENT, "column index out of range");
    if (r->streamed) return fail(QE_ERR_STATE, "streamed result: the rows are in the host buffers given to qe_pipeline_stream");
    qe_ctx *ctx = r->ctx;
    DeviceGuard g(ctx->device);
    const ResultColumn &c = r->cols[k];
    if (r->mode == qe::KM_AGG) {
        double v = 0.0;
        int is_null = 0;
        rc = qe_result_agg(r, k, &v, &is_null);
        if (rc) return rc;
        if (dst_values) *(double *)dst_values = v;
        if (dst_validity) *dst_validity = is_null ? 0 : 1;
        return QE_OK;
    }
    if (n == 0) return QE_OK;
    if (is_groupby(r)) {                       // finished on the host by qe_groupby_run
        if (dst_values) memcpy(dst_values, r->h_values[k].data(), r->h_values[k].size());
        if (dst_validity) memcpy(dst_validity, r->h_valid[k].data(), r->h_valid[k].size());
        return QE_OK;
    }
    if (dst_values) {
        const size_t bytes = c.type == QE_TYPE_BOOLEAN ? (size_t)((n + 7) / 8) : (size_t)n * type_width(c.type);
        QE_CUDA(cudaMemcpyAsync(dst_values, c.values.ptr, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    }
    if (dst_validity) {
        if (c.has_validity)
            QE_CUDA(cudaMemcpyAsync(dst_validity, c.valid.ptr, (size_t)((n + 7) / 8), cudaMemcpyDeviceToHost, ctx->stream));
        else
            fill_valid_bitmap(dst_validity, n);
    }
    QE_CUDA(cudaStreamSynchronize(ctx->stream));
    return QE_OK;
}

extern "C" int qe_result_selection(qe_result *r, uint32_t *dst)
{
    int64_t n = 0;
    int rc = qe_result_num_rows(r, &n);
    if (rc) return rc;
    if (r->mode != qe::KM_FILTER || !r->sel.ptr) return fail(QE_ERR_STATE, "this result has no selection vector");
    if (r->streamed) return fail(QE_ERR_STATE, "streamed result: the selection vector is in the host buffer given to qe_pipeline_stream");
    if (!dst) return fail(QE_ERR_INVALID_ARGUMENT, "dst is NULL");
    if (n == 0) return QE_OK;
    DeviceGuard g(r->ctx->device);
    QE_CUDA(cudaMemcpyAsync(dst, r->sel.ptr, (size_t)n * 4, cudaMemcpyDeviceToHost, r->ctx->stream));
    QE_CUDA(cudaStreamSynchronize(r->ctx->stream));
    return QE_OK;
}

static qe_agg_partial partial_of(const qe_result *r, int k)
{
    qe_agg_partial p;
    p.sum = r->host->agg[k][0];
    memcpy(&p.count, &r->host->agg[k][1], sizeof p.count);
    p.min = r->host->agg[k][2];
    p.max = r->host->agg[k][3];
    return p;
}

extern "C" int qe_result_agg_partial(qe_result *r, int k, qe_agg_partial *out)
{
    int rc = wait_result(r);
    if (rc) return rc;
    if (r->mode != qe::KM_AGG) return fail(QE_ERR_STATE, "this result is not an aggregation");
    if (k < 0 || k >= (int)r->agg_fns.size() || !out) return fail(QE_ERR_INVALID_ARGUMENT, "aggregate index out of range");
    *out = partial_of(r, k);
    if (r->batch_rows == 0) {
        out->count = 0;
        out->sum = 0.0;
    }
    return QE_OK;
}

extern "C" int qe_result_agg_partials(qe_result *r, int k, qe_agg_partial *dst, int64_t capacity, int64_t *n)
{
    int rc = wait_result(r);
    if (rc) return rc;
    if (r->mode != qe::KM_AGG) return fail(QE_ERR_STATE, "this result is not an aggregation");
    const size_t nagg = r->agg_fns.size();
    if (k < 0 || k >= (int)nagg) return fail(QE_ERR_INVALID_ARGUMENT, "aggregate index out of range");
    const int64_t have = r->batch_rows == 0 ? 0 : r->n_super;
    if (n) *n = have;
    for (int64_t s = 0; dst && s < have && s < capacity; s++) {
        const double *src = r->h_super + ((size_t)s * nagg + (size_t)k) * 4;
        dst[s].sum = src[0];
        memcpy(&dst[s].count, &src[1], sizeof dst[s].count);
        dst[s].min = src[2];
        dst[s].max = src[3];
    }
    return QE_OK;
}

// finish() of the reference's accumulators, M/evaluator/Accumulators.kt:33-35,47-53,65-71,83-89,101-107.
// COUNT is int64 here; the reference's Kotlin Int would wrap above 2^31-1 rows.
static void finish_agg(int fn, const qe_agg_partial &p, double *value, int *is_null)
{
    double v = 0.0;
    int null = 0;
    switch (fn) {
    case QE_AGG_FN_COUNT: v = (double)p.count; break;
    case QE_AGG_FN_SUM: if (p.count == 0) null = 1; else v = p.sum; break;
    case QE_AGG_FN_AVG: if (p.count == 0) null = 1; else v = p.sum / (double)p.count; break;
    case QE_AGG_FN_MIN: if (p.count == 0) null = 1; else v = p.min; break;
    case QE_AGG_FN_MAX: if (p.count == 0) null = 1; else v = p.max; break;
    }
    if (value) *value = v;
    if (is_null) *is_null = null;
}

extern "C" int qe_result_agg(qe_result *r, int k, double *value, int *is_null)
{
    qe_agg_partial p;
    int rc = qe_result_agg_partial(r, k, &p);
    if (rc) return rc;
    finish_agg(r->agg_fns[k], p, value, is_null);
    return QE_OK;
}

extern "C" int qe_agg_combine(int agg_fn, const qe_agg_partial *partials, int n, double *value, int *is_null)
{
    if (agg_fn < 0 || agg_fn > 4) return fail(QE_ERR_INVALID_ARGUMENT, "unknown aggregation function");
    if (n < 0 || (n > 0 && !partials)) return fail(QE_ERR_INVALID_ARGUMENT, "bad partial list");
    qe_agg_partial t;
    t.sum = 0.0;
    t.count = 0;
    t.min = qe_bits_f64((qe_i64)0x7ff0000000000000LL);
    t.max = qe_bits_f64((qe_i64)0xfff0000000000000ULL);
    for (int i = 0; i < n; i++) {          // array order = rank order: deterministic
        if (partials[i].count == 0) continue;
        t.count += partials[i].count;
        t.sum += partials[i].sum;
        t.min = qe_math_min(t.min, partials[i].min);
        t.max = qe_math_max(t.max, partials[i].max);
    }
    finish_agg(agg_fn, t, value, is_null);
    return QE_OK;
}

extern "C" void *qe_result_device_ptr(qe_result *r, int k, int which)
{
    if (!r || r->streamed) return nullptr;
    if (which == 2) return r->sel.ptr;
    if (k < 0 || k >= (int)r->cols.size()) return nullptr;
    return which == 0 ? r->cols[k].values.ptr : r->cols[k].valid.ptr;
}

extern "C" int qe_result_kernel_ms(qe_result *r, float *ms)
{
    int rc = wait_result(r);
    if (rc) return rc;
    if (!ms) return fail(QE_ERR_INVALID_ARGUMENT, "ms is NULL");
    DeviceGuard g(r->ctx->device);
    QE_CUDA(cudaEventElapsedTime(ms, r->ev_begin, r->ev_end));     // events keep their timestamps after folding
    return QE_OK;
}

extern "C" int qe_result_kernel_stats(qe_result *r, double *total_ms, int64_t *runs)
{
    int rc = wait_result(r);
    if (rc) return rc;
    DeviceGuard g(r->ctx->device);
    fold_timing(r);
    double ms = r->total_kernel_ms;
    int64_t nr = r->kernel_runs;
    for (qe_result *m : r->ring) {                             // bounded streaming: the other output buffer sets ran chunks too
        fold_timing(m);
        ms += m->total_kernel_ms;
        nr += m->kernel_runs;
    }
    if (total_ms) *total_ms = ms;
    if (runs) *runs = nr;
    return QE_OK;
}

extern "C" int qe_result_debug_counters(qe_result *r, uint64_t *dst, size_t n)
{
    int rc = wait_result(r);
    if (rc) return rc;
    if (!r->debug.ptr) return fail(QE_ERR_STATE, "no debug counters were recorded");
    DeviceGuard g(r->ctx->device);
    QE_CUDA(cudaMemcpy(dst, r->debug.ptr, std::min(n, (size_t)4096 * 8) * sizeof(uint64_t), cudaMemcpyDeviceToHost));
    return QE_OK;
}

extern "C" int qe_result_used_jit(const qe_result *r) { return r && r->used_jit ? 1 : 0; }

// --------------------------------------------------------------- streaming
// Host columns in, host results out, in row-range chunks: the upload of chunk c+1 (copy engine, h2d
// stream), the kernels of chunk c (compute stream) and the download of the rows chunk c-1 produced
// (other copy engine, d2h stream) run at the same time, so a scan that starts and ends in host memory
// is bound by the busier PCIe direction instead of by the sum of both plus the kernels.
//   resident mode: every chunk after the first runs with QE_RUN_APPEND, the device-side results are the
//     same as one run over the whole table.
//   bounded mode: the table is a ring of chunk slots and every pipeline has as many output buffer sets;
//     a chunk's kernel starts its output positions at (rows kept so far) & 7, so compacted bit columns keep
//     their phase and the chunks' bytes are laid end to end in the caller's buffers (the one byte two chunks
//     share goes through a scratch byte and is OR-ed in at the end).

namespace {

struct ByteFix { uint8_t *dst; size_t scratch; };

struct StreamState {
    qe_ctx *ctx = nullptr;
    std::vector<cudaEvent_t> events;
    unsigned long long *counts = nullptr;      // pinned: rows kept up to and including chunk c, per pipeline
    uint8_t *scratch = nullptr;                // pinned: first bytes of bit columns that share a byte with the chunk before
    size_t scratch_used = 0, scratch_cap = 0;
    std::vector<ByteFix> fixes;
    ~StreamState()
    {
        if (ctx) {
            if (ctx->h2d) cudaStreamSynchronize(ctx->h2d);
            cudaStreamSynchronize(ctx->stream);
            if (ctx->d2h) cudaStreamSynchronize(ctx->d2h);
        }
        for (cudaEvent_t e : events) if (e) cudaEventDestroy(e);
        if (counts) cudaFreeHost(counts);
        if (scratch) cudaFreeHost(scratch);
    }
    cudaError_t new_event(cudaEvent_t *e)
    {
        cudaError_t rc = cudaEventCreateWithFlags(e, cudaEventDisableTiming);
        if (rc == cudaSuccess) events.push_back(*e);
        return rc;
    }
};

// bits [from, to) of a host bitmap <- bits [src_from, ...) of a device bitmap, src_from == from (mod 8).
// `own_first_byte`: the device byte holding bit src_from has no bits of earlier chunks (bounded mode), so it must be
// OR-ed into the host byte instead of replacing it.
int download_bits(StreamState &st, uint8_t *host, const void *dev, int64_t from, int64_t to, int64_t src_from, bool own_first_byte)
{
    if (to <= from) return QE_OK;
    size_t hb = (size_t)(from / 8), db = (size_t)(src_from / 8);
    size_t nb = (size_t)((to + 7) / 8) - hb;
    if (own_first_byte && (from % 8)) {
        if (st.scratch_used >= st.scratch_cap) return fail(QE_ERR_STATE, "stream scratch exhausted");
        QE_CUDA(cudaMemcpyAsync(st.scratch + st.scratch_used, (const char *)dev + db, 1, cudaMemcpyDeviceToHost, st.ctx->d2h));
        st.fixes.push_back(ByteFix{host + hb, st.scratch_used++});
        hb++; db++; nb--;
    }
    if (nb) QE_CUDA(cudaMemcpyAsync(host + hb, (const char *)dev + db, nb, cudaMemcpyDeviceToHost, st.ctx->d2h));
    return QE_OK;
}

// output rows [from, to) of pipeline q, which sit at position src_from of `r`'s device buffers, to the caller's buffers
int download_range(StreamState &st, const qe_result *r, int q, const qe_stream_desc *sd, int64_t from, int64_t to, int64_t src_from,
                   bool own_first_byte)
{
    if (to <= from) return QE_OK;
    qe_ctx *ctx = st.ctx;
    if (sd->host_selection && sd->host_selection[q] && r->sel.ptr)
        QE_CUDA(cudaMemcpyAsync(sd->host_selection[q] + from, (const uint32_t *)r->sel.ptr + src_from, (size_t)(to - from) * 4,
                                cudaMemcpyDeviceToHost, ctx->d2h));
    for (size_t k = 0; k < r->cols.size(); k++) {
        const ResultColumn &c = r->cols[k];
        void *hv = sd->host_out_values && sd->host_out_values[q] ? sd->host_out_values[q][k] : nullptr;
        uint8_t *hn = sd->host_out_validity && sd->host_out_validity[q] ? sd->host_out_validity[q][k] : nullptr;
        int rc = QE_OK;
        if (hv) {
            if (c.type == QE_TYPE_BOOLEAN)
                rc = download_bits(st, (uint8_t *)hv, c.values.ptr, from, to, src_from, own_first_byte);
            else
                QE_CUDA(cudaMemcpyAsync((char *)hv + (size_t)from * type_width(c.type),
                                        (const char *)c.values.ptr + (size_t)src_from * type_width(c.type),
                                        (size_t)(to - from) * type_width(c.type), cudaMemcpyDeviceToHost, ctx->d2h));
            if (rc) return rc;
        }
        if (hn) {
            if (c.has_validity) {
                if ((rc = download_bits(st, hn, c.valid.ptr, from, to, src_from, own_first_byte))) return rc;
            } else {                                                            // statically non-null: all ones
                for (int64_t i = from; i < to && (i % 8); i++) hn[i / 8] |= (uint8_t)(1u << (i % 8));
                const int64_t a = (from + 7) / 8 * 8;
                if (to > a) {
                    memset(hn + a / 8, 0xff, (size_t)((to - a) / 8));
                    if ((to - a) % 8) hn[to / 8] = (uint8_t)((1u << (to % 8)) - 1u);
                }
            }
        }
    }
    return QE_OK;
}

}  // namespace

extern "C" int qe_pipeline_stream(qe_ctx *ctx, qe_table *table, const qe_stream_desc *sd)
{
    if (!ctx || !table || !sd || !sd->pipelines || !sd->results || sd->npipelines <= 0)
        return fail(QE_ERR_INVALID_ARGUMENT, "NULL argument");
    if (table->ctx != ctx) return fail(QE_ERR_INVALID_ARGUMENT, "table belongs to another context");
    if (sd->npipelines > 16) return fail(QE_ERR_UNSUPPORTED, "at most 16 pipelines per stream call");
    const int64_t chunk = sd->chunk_rows > 0 ? sd->chunk_rows : (int64_t)32 << 20;
    if (chunk % 128) return fail(QE_ERR_INVALID_ARGUMENT, "chunk_rows must be a multiple of 128");
    if (sd->total_rows < 0) return fail(QE_ERR_INVALID_ARGUMENT, "total_rows must not be negative");
    const bool bounded = sd->total_rows > 0;
    const bool relative = (sd->flags & QE_STREAM_CHUNK_RELATIVE) != 0;
    const int np = sd->npipelines;
    const int64_t n = bounded ? sd->total_rows : table->nrows;
    int nslots = 1;
    if (bounded) {
        const int64_t fit = table->nrows / chunk;
        if (fit < 2) return fail(QE_ERR_INVALID_ARGUMENT, "bounded streaming: the table must hold at least two chunks (nrows >= 2 * chunk_rows)");
        // Every slot is a chunk of the table plus one output buffer set per pipeline.  Three keep upload, kernels and
        // download of neighbouring chunks busy; more let the uploads run ahead of the downloads when those are the slower
        // direction (the window the caller allocates decides).
        nslots = (int)std::min<int64_t>(fit, 32);
    }
    if (chunk > 0xffffffffLL) return fail(QE_ERR_INVALID_ARGUMENT, "chunk_rows must fit 32 bits");
    const int64_t nchunks = std::max<int64_t>(1, (n + chunk - 1) / chunk);
    if (nchunks > (1 << 20)) return fail(QE_ERR_INVALID_ARGUMENT, "more than 2^20 chunks: use a larger chunk_rows");
    for (int q = 0; q < np; q++) {
        const qe_pipeline_desc &d = sd->pipelines[q];
        if (n > 0x100000000LL && !relative && d.filter && d.nagg == 0 && !(d.flags & QE_RUN_NO_SELECTION))
            return fail(QE_ERR_INVALID_ARGUMENT, "more than 2^32 rows: selection vectors need QE_STREAM_CHUNK_RELATIVE");
    }
    const int64_t spc = (chunk + QE_AGG_SUPER_ROWS - 1) / QE_AGG_SUPER_ROWS;       // super-tile partials per chunk
    DeviceGuard g(ctx->device);
    if (!ctx->h2d) QE_CUDA(cudaStreamCreateWithFlags(&ctx->h2d, cudaStreamNonBlocking));
    if (!ctx->d2h) QE_CUDA(cudaStreamCreateWithFlags(&ctx->d2h, cudaStreamNonBlocking));

    StreamState st;
    st.ctx = ctx;
    QE_CUDA(cudaHostAlloc((void **)&st.counts, sizeof(unsigned long long) * (size_t)nchunks * np, cudaHostAllocDefault));
    st.scratch_cap = bounded ? (size_t)nchunks * (size_t)np * 2 * QE_MAX_PROJECTIONS + 16 : 16;
    QE_CUDA(cudaHostAlloc((void **)&st.scratch, st.scratch_cap, cudaHostAllocDefault));
    std::vector<cudaEvent_t> ev_up((size_t)nchunks, nullptr), ev_run((size_t)nchunks, nullptr), ev_dl((size_t)nchunks, nullptr);
    // bounded mode: output buffer sets of pipeline q, [0] = results[q] itself, the others are owned by it
    std::vector<std::vector<qe_result *>> ring((size_t)np, std::vector<qe_result *>((size_t)nslots, nullptr));
    for (int q = 0; q < np; q++) {
        qe_result *r = sd->results[q];
        if (r && r->ctx != ctx) return fail(QE_ERR_INVALID_ARGUMENT, "result belongs to another context");
        if (r && r->pending) { QE_CUDA(cudaEventSynchronize(r->ev_done)); r->pending = false; }
        for (int j = 1; r && j < nslots && (size_t)(j - 1) < r->ring.size(); j++) ring[(size_t)q][(size_t)j] = r->ring[(size_t)(j - 1)];
    }
    auto slot_result = [&](int q, int64_t c) -> qe_result *& {
        const size_t j = bounded ? (size_t)(c % nslots) : 0;
        return j == 0 ? sd->results[q] : ring[(size_t)q][j];
    };

    // earlier work on the compute stream (generators, uploads through qe_table_upload) comes first
    cudaEvent_t ev_start = nullptr;
    QE_CUDA(st.new_event(&ev_start));
    QE_CUDA(cudaEventRecord(ev_start, ctx->stream));
    QE_CUDA(cudaStreamWaitEvent(ctx->h2d, ev_start, 0));

    int rc = QE_OK;
    auto drain = [&](int64_t c) -> int {                       // download what chunk c added, pipeline by pipeline
        QE_CUDA(cudaEventSynchronize(ev_run[(size_t)c]));
        const int64_t r0 = c * chunk, r1 = std::min(n, r0 + chunk);
        for (int q = 0; q < np; q++) {
            const qe_pipeline_desc &d = sd->pipelines[q];
            if (d.nagg > 0) continue;                           // nothing to download per chunk
            const qe_result *src = slot_result(q, c);
            const bool dense = src->mode == qe::KM_PROJECT;
            const int64_t kb = dense ? r0 : (c > 0 ? (int64_t)st.counts[(size_t)(c - 1) * np + q] : 0);
            int64_t ka = dense ? r1 : (int64_t)st.counts[(size_t)c * np + q];
            if (sd->chunk_kept && sd->chunk_kept[q]) sd->chunk_kept[q][c] = ka;
            if (bounded) {
                const int64_t cap = d.out_capacity_rows > 0 ? d.out_capacity_rows : n;
                if (ka > cap) return fail(QE_ERR_CAPACITY, "filter kept " + std::to_string(ka) + " rows but out_capacity_rows is " + std::to_string(cap));
            } else {
                ka = std::min(ka, src->capacity);               // qe_result_num_rows reports the overflow
            }
            const int drc = download_range(st, src, q, sd, kb, ka, bounded ? (dense ? 0 : (kb & 7)) : kb, bounded && !dense);
            if (drc) return drc;
        }
        QE_CUDA(st.new_event(&ev_dl[(size_t)c]));
        QE_CUDA(cudaEventRecord(ev_dl[(size_t)c], ctx->d2h));
        return QE_OK;
    };

    for (int64_t c = 0; c < nchunks; c++) {
        const int64_t r0 = c * chunk, r1 = std::min(n, r0 + chunk), nc = r1 - r0;
        const int64_t base_row = bounded ? (c % nslots) * chunk : r0;          // where the chunk lives in `table`
        // ---- upload (or generate) chunk c
        if (bounded && c >= nslots) QE_CUDA(cudaStreamWaitEvent(ctx->h2d, ev_run[(size_t)(c - nslots)], 0));   // the slot's last reader
        for (size_t col = 0; col < table->cols.size() && nc > 0; col++) {
            const qe_col_desc &cd = table->cols[col];
            const void *hv = sd->host_values ? sd->host_values[col] : nullptr;
            const uint8_t *hn = sd->host_validity ? sd->host_validity[col] : nullptr;
            const size_t w = type_width(cd.type);
            if (sd->gen && sd->gen[col].kind >= 0) {
                const qe_stream_gen &gs = sd->gen[col];
                if (gs.kind > 4) return fail(QE_ERR_INVALID_ARGUMENT, "unknown generator kind");
                void *dv = cd.type == QE_TYPE_BOOLEAN ? (void *)((char *)table->values[col] + base_row / 8)
                                                      : (void *)((char *)table->values[col] + (size_t)base_row * w);
                QE_CUDA(qe::launch_generate(dv, table->valid[col] ? table->valid[col] + base_row / 32 : nullptr, cd.type, gs.kind, gs.seed,
                                            gs.p0, gs.p1, cd.nullable ? gs.null_rate : 0.0, sd->gen_row_offset + r0, nc, ctx->stream));
                ctx->launches++;
                continue;
            }
            if (hv) {
                if (cd.type == QE_TYPE_BOOLEAN)
                    QE_CUDA(cudaMemcpyAsync((char *)table->values[col] + base_row / 8, (const char *)hv + r0 / 8,
                                            (size_t)((nc + 7) / 8), cudaMemcpyHostToDevice, ctx->h2d));
                else
                    QE_CUDA(cudaMemcpyAsync((char *)table->values[col] + (size_t)base_row * w, (const char *)hv + (size_t)r0 * w,
                                            (size_t)nc * w, cudaMemcpyHostToDevice, ctx->h2d));
            }
            if (hn && cd.nullable)
                QE_CUDA(cudaMemcpyAsync((char *)table->valid[col] + base_row / 8, hn + r0 / 8, (size_t)((nc + 7) / 8),
                                        cudaMemcpyHostToDevice, ctx->h2d));
        }
        QE_CUDA(st.new_event(&ev_up[(size_t)c]));
        QE_CUDA(cudaEventRecord(ev_up[(size_t)c], ctx->h2d));
        QE_CUDA(cudaStreamWaitEvent(ctx->stream, ev_up[(size_t)c], 0));
        // the output buffer set this chunk writes must have been downloaded
        if (bounded && c >= nslots) QE_CUDA(cudaStreamWaitEvent(ctx->stream, ev_dl[(size_t)(c - nslots)], 0));
        // ---- run every pipeline over chunk c
        for (int q = 0; q < np; q++) {
            qe_pipeline_desc d = sd->pipelines[q];
            d.row_begin = base_row;
            d.row_end = base_row + nc;
            d.sel_offset = (relative ? 0 : r0) - base_row;
            d.flags &= ~(uint32_t)QE_RUN_APPEND;
            const bool is_agg = d.nagg > 0;
            RunExtra extra;
            qe_result **slot = &sd->results[q];
            if (is_agg) {
                extra.super_offset = c * spc;
                extra.super_total = nchunks * spc;
            } else if (bounded) {
                slot = &slot_result(q, c);
                extra.chained = true;
                extra.chain_base = c > 0 ? (const qe_u64 *)slot_result(q, c - 1)->ctl.ptr : nullptr;
                d.out_capacity_rows = nc + 8;
            } else {
                if (c > 0) d.flags |= (uint32_t)QE_RUN_APPEND;
                if (d.out_capacity_rows <= 0) d.out_capacity_rows = n;
            }
            const bool fresh = *slot == nullptr;
            if ((rc = run_pipeline(ctx, table, &d, slot, &extra))) return rc;
            qe_result *r = *slot;
            if (fresh && slot != &sd->results[q]) sd->results[q]->ring.push_back(r);      // owned by the pipeline's result
            if (bounded)
                QE_CUDA(qe::launch_store_u64(&st.counts[(size_t)c * np + q], (const unsigned long long *)r->ctl.ptr, ctx->stream));
            else
                QE_CUDA(cudaMemcpyAsync(&st.counts[(size_t)c * np + q], r->ctl.ptr, sizeof(unsigned long long),
                                        cudaMemcpyDeviceToHost, ctx->stream));
        }
        QE_CUDA(st.new_event(&ev_run[(size_t)c]));
        QE_CUDA(cudaEventRecord(ev_run[(size_t)c], ctx->stream));
        // ---- while chunk c runs, download what chunk c-1 produced
        if (c > 0 && (rc = drain(c - 1))) return rc;
    }
    if ((rc = drain(nchunks - 1))) return rc;
    QE_CUDA(cudaStreamSynchronize(ctx->h2d));
    QE_CUDA(cudaStreamSynchronize(ctx->stream));
    QE_CUDA(cudaStreamSynchronize(ctx->d2h));
    for (const ByteFix &f : st.fixes) *f.dst |= st.scratch[f.scratch];
    for (int q = 0; q < np; q++) {                             // the runs are complete: make the results readable
        qe_result *r = sd->results[q];
        r->pending = false;
        for (qe_result *m : r->ring) m->pending = false;
        if (r->mode == qe::KM_AGG) {
            r->batch_rows = n;                                  // n_super: the last chunk's run left the length of the whole list
            r->agg_folded = false;
        } else if (bounded) {
            r->streamed = true;
            r->batch_rows = n;
            r->host->count = r->mode == qe::KM_PROJECT ? (unsigned long long)n : st.counts[(size_t)(nchunks - 1) * np + q];
            // the next stream of the same query picks its kernel shape by what this one kept (every chunk the same shape:
            // the hint only changes here, between streams)
            if (r->mode == qe::KM_FILTER && n > 0) {
                r->f_sel = (double)r->host->count / (double)n;
                for (qe_result *m : r->ring) m->f_sel = r->f_sel;
                if (!r->f_sig.empty()) ctx->sel_memo[r->f_sig] = r->f_sel;
            }
        }
    }
    return QE_OK;
}

// ------------------------------------------------------------ diagnostics
// Bind a pipeline against a schema, generate its CUDA source and optionally
// compile it with NVRTC.  Needs no GPU: used by the CPU-side tests and for
// inspecting the generated code (SASS via cuobjdump on the returned cubin).
struct DebugOpts {
    int steps = 0, vec = 0, staging = -1;
    double sel_hint = -1.0;       // filter: selectivity the same query measured last time
    int min_blocks = 0;           // dense projection: 4 = the tight variant tried first
    int seen_groups = 0;          // group-by, register variant: groups found last time
    uint32_t run_flags = 0;       // QE_RUN_NO_VALUES
    bool runtime_shapes = false;  // group-by: the shape qe_groupby_run would launch (register variant only)
};

static int debug_generate(const qe_col_desc *cols, int ncols, const uint32_t *const *programs, const size_t *nwords,
                          int has_filter, int nproj, const int32_t *agg_fns, int nagg, int group_keys, int atomic,
                          const DebugOpts &o, int compile,
                          char *src_buf, size_t src_cap, char *cubin_buf, size_t cubin_cap, size_t *cubin_bytes,
                          char *key_out = nullptr, int *origin_out = nullptr)
{
    if (!cols || ncols <= 0 || !programs || !nwords) return fail(QE_ERR_INVALID_ARGUMENT, "NULL argument");
    std::vector<qe::ColumnInfo> schema(ncols);
    for (int c = 0; c < ncols; c++) { schema[c].type = cols[c].type; schema[c].nullable = cols[c].nullable != 0; }
    const int nprog = (has_filter ? 1 : 0) + nproj;
    std::vector<qe::Program> progs(nprog);
    std::string err;
    for (int i = 0; i < nprog; i++)
        if (!qe::parse_program(programs[i], nwords[i], progs[i], err)) return fail(QE_ERR_PROGRAM, err);
    const bool no_values = (o.run_flags & QE_RUN_NO_VALUES) && has_filter && nagg == 0 && group_keys < 0;
    std::vector<const qe::Program *> proj;
    for (int k = 0; k < nproj && !no_values; k++) proj.push_back(&progs[(has_filter ? 1 : 0) + k]);
    std::vector<int> aggs;
    if (nagg > 0) aggs.assign(agg_fns, agg_fns + nagg);
    qe::BoundPipeline bp;
    int rc = qe::bind_pipeline(schema, has_filter ? &progs[0] : nullptr, proj, aggs, bp, err, group_keys);
    if (rc) return fail(rc, err);
    const qe::KernelMode mode = group_keys >= 0 ? (atomic == 1 ? qe::KM_GROUPBY_ATOMIC : (atomic == 2 || (atomic >= 4 && atomic <= 11)) ? qe::KM_GROUPBY_SHARED : atomic == 3 ? qe::KM_GROUPBY_REG : qe::KM_GROUPBY_PRIVATE)
                              : !bp.agg_fns.empty() ? qe::KM_AGG : bp.has_filter ? qe::KM_FILTER : qe::KM_PROJECT;
    qe_ctx dummy;
    dummy.opt_steps = o.steps;
    dummy.opt_vec = o.vec;
    dummy.opt_staging = o.staging;
    qe::TileShape shape = choose_shape(&dummy, bp, mode, o.sel_hint);
    if (o.min_blocks > 0) shape.min_blocks = o.min_blocks;
    if (mode == qe::KM_GROUPBY_SHARED) { shape.gthreads = 512; shape.steps = 1; shape.vec = qe::jit_wide_loads_supported() ? 4 : 2; shape.scopies = atomic == 4 ? 32 : atomic == 5 ? 8 : atomic == 7 ? 16 : atomic == 8 ? 4 : atomic == 9 ? 2 : 1; }
    if (mode == qe::KM_GROUPBY_SHARED && atomic == 6) { shape = choose_shape(&dummy, bp, mode); shape.gthreads = 256; shape.scopies = -1; }
    if (mode == qe::KM_GROUPBY_SHARED && (atomic == 10 || atomic == 11)) {       // the two passes (scatter, gather with 8 copies)
        const qe::GroupPlan g2 = qe::plan_groupby(bp, 512);
        qe::GroupPlan::scatter_shape(g2.rec_words, shape.npart, shape.pcap);
        shape.gpass = atomic == 10 ? 1 : 2;
        shape.scopies = 8;
    }
    if (mode == qe::KM_GROUPBY_REG && o.runtime_shapes) {       // as qe_groupby_run, level 0
        const qe::GroupPlan gp = qe::plan_groupby(bp, 256);
        if (gp.rmax < 2 || !shape.greg) return fail(QE_ERR_UNSUPPORTED, "the register variant does not apply to this group-by");
        shape.rmax_hint = o.seen_groups > 0 ? std::min(gp.rmax, std::max(4, o.seen_groups)) : gp.rmax;
        shape.scopies = 1;
    }
    const std::string src = qe::generate_source(bp, mode, shape);
    if (src_buf && src_cap) {
        const size_t n = std::min(src_cap - 1, src.size());
        memcpy(src_buf, src.data(), n);
        src_buf[n] = 0;
    }
    if (cubin_bytes) *cubin_bytes = 0;
    if (key_out) {
        const std::string key = qe::cubin_key(src, std::vector<std::string>());
        memcpy(key_out, key.c_str(), std::min<size_t>(key.size() + 1, 33));
    }
    if (compile) {
        qe::CubinRef cubin;
        std::string log;
        int origin = 0;
        const int crc = qe::cubin_get(src, std::vector<std::string>(), cubin, &origin, log);
        if (crc) return fail(QE_ERR_JIT, log);
        if (origin_out) *origin_out = origin;
        if (cubin_bytes) *cubin_bytes = cubin->size();
        if (cubin_buf && cubin_cap >= cubin->size()) memcpy(cubin_buf, cubin->data(), cubin->size());
    }
    return QE_OK;
}

extern "C" int qe_debug_generate(const qe_col_desc *cols, int ncols,
                                 const uint32_t *const *programs, const size_t *nwords,
                                 int has_filter, int nproj, const int32_t *agg_fns, int nagg,
                                 int steps, int vec, int staging, int compile,
                                 char *src_buf, size_t src_cap, char *cubin_buf, size_t cubin_cap, size_t *cubin_bytes)
{
    DebugOpts o;
    o.steps = steps;
    o.vec = vec;
    o.staging = staging;
    return debug_generate(cols, ncols, programs, nwords, has_filter, nproj, agg_fns, nagg, -1, 0, o, compile,
                          src_buf, src_cap, cubin_buf, cubin_cap, cubin_bytes);
}

extern "C" int qe_debug_generate_groupby(const qe_col_desc *cols, int ncols,
                                         const uint32_t *const *programs, const size_t *nwords,
                                         int has_filter, int nkeys, const int32_t *agg_fns, int nagg, int atomic, int compile,
                                         char *src_buf, size_t src_cap, size_t *cubin_bytes)
{
    if (nkeys < 0 || nagg < 0) return fail(QE_ERR_INVALID_ARGUMENT, "bad key / aggregate count");
    return debug_generate(cols, ncols, programs, nwords, has_filter, nkeys + nagg, agg_fns, nagg, nkeys, atomic, DebugOpts(), compile,
                          src_buf, src_cap, nullptr, 0, cubin_bytes);
}

// Ahead-of-time compilation (no GPU needed): the kernel qe_pipeline_run_ex / qe_groupby_run would ask for, with the
// context options at their defaults, is generated and compiled through the cubin caches -- i.e. it lands in
// $QE_CACHE_DIR, from where queryengine_b200/build.py bakes it into the library (qe_aot_table.cpp).
extern "C" int qe_aot_compile(const qe_aot_desc *d, char key_out[33], int *origin)
{
    if (!d) return fail(QE_ERR_INVALID_ARGUMENT, "desc is NULL");
    DebugOpts o;
    o.sel_hint = d->sel_hint;
    o.min_blocks = d->min_blocks;
    o.seen_groups = d->seen_groups;
    o.run_flags = d->run_flags;
    o.runtime_shapes = true;
    const bool group = d->group_keys >= 0;
    return debug_generate(d->cols, d->ncols, d->programs, d->nwords, d->has_filter, d->nproj, d->agg_fns, d->nagg,
                          group ? d->group_keys : -1, group ? 3 : 0, o, 1, nullptr, 0, nullptr, 0, nullptr, key_out, origin);
}
