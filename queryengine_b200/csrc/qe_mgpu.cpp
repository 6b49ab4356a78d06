// qe_mgpu.cpp -- several GPUs driven by one process, on top of the single-device C ABI (include/qegpu.h,
// "several GPUs, one process").  Host code only: one context and one worker thread per device, every call fans
// out to all devices and joins; partial aggregates and group-by partials are combined on the calling thread in
// partition order.  Nothing here evaluates a row.
#include <algorithm>
#include <condition_variable>
#include <cstring>
#include <deque>
#include <functional>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

#include "../../include/qegpu.h"
#include "qe_internal.h"

namespace {

struct Worker {
    std::thread th;
    std::mutex mu;
    std::condition_variable cv;
    std::deque<std::function<void()>> q;
    bool stop = false;

    Worker() { th = std::thread([this] { loop(); }); }
    ~Worker()
    {
        {
            std::lock_guard<std::mutex> g(mu);
            stop = true;
        }
        cv.notify_all();
        if (th.joinable()) th.join();
    }
    void loop()
    {
        for (;;) {
            std::function<void()> job;
            {
                std::unique_lock<std::mutex> g(mu);
                cv.wait(g, [this] { return stop || !q.empty(); });
                if (q.empty()) return;
                job = std::move(q.front());
                q.pop_front();
            }
            job();
        }
    }
    void post(std::function<void()> job)
    {
        {
            std::lock_guard<std::mutex> g(mu);
            q.push_back(std::move(job));
        }
        cv.notify_one();
    }
};

size_t width_of(int type)
{
    switch (type) {
    case QE_TYPE_DOUBLE: case QE_TYPE_INT64: return 8;
    case QE_TYPE_DATE32: case QE_TYPE_STRING: return 4;
    default: return 0;      // BOOLEAN: bitmap
    }
}

// bits [0, n) of src appended to dst at bit position `at`; dst is zero from `at` on and holds ceil((at + n) / 8) bytes
void append_bits(uint8_t *dst, int64_t at, const uint8_t *src, int64_t n)
{
    if (n <= 0) return;
    const int sh = (int)(at & 7);
    uint8_t *d = dst + at / 8;
    const int64_t full = n / 8;
    const int rem = (int)(n & 7);
    for (int64_t i = 0; i < full + (rem ? 1 : 0); i++) {
        uint8_t b = src[i];
        if (i == full) b &= (uint8_t)((1u << rem) - 1u);         // bits past the end of src do not count
        d[i] |= (uint8_t)(b << sh);
        if (sh && (b >> (8 - sh))) d[i + 1] |= (uint8_t)(b >> (8 - sh));
    }
}

}  // namespace

struct qe_mgpu {
    std::vector<int> devices;
    std::vector<qe_ctx *> ctx;
    std::vector<std::unique_ptr<Worker>> workers;
};

struct qe_mtable {
    qe_mgpu *m = nullptr;
    int64_t nrows = 0;
    std::vector<qe_col_desc> cols;
    std::vector<qe_table *> parts;
    std::vector<int64_t> begin, end;
};

struct qe_mresult {
    qe_mgpu *m = nullptr;
    std::vector<qe_result *> parts;
    std::vector<int64_t> begin;                 // first row of every partition
    std::vector<int> agg_fns;                   // global aggregation: the functions
    bool is_agg = false, is_group = false;
    // group-by: the merged groups, finished on the host
    int g_nkeys = 0;
    int64_t g_rows = 0;
    std::vector<int> g_types;
    std::vector<std::vector<uint8_t>> g_values, g_valid;
};

// fn(part) on every device's worker thread at once; the first failure (by partition) is reported on this thread
template <class F>
static int fan_out(qe_mgpu *m, F fn)
{
    const int n = (int)m->ctx.size();
    std::vector<int> rc((size_t)n, 0);
    std::vector<std::string> msg((size_t)n);
    std::mutex mu;
    std::condition_variable cv;
    int left = n;
    for (int p = 0; p < n; p++) {
        m->workers[(size_t)p]->post([&, p] {
            const int r = fn(p);
            if (r) msg[(size_t)p] = qe_last_error();
            std::lock_guard<std::mutex> g(mu);
            rc[(size_t)p] = r;
            if (--left == 0) cv.notify_all();
        });
    }
    {
        std::unique_lock<std::mutex> g(mu);
        cv.wait(g, [&] { return left == 0; });
    }
    for (int p = 0; p < n; p++)
        if (rc[(size_t)p]) return qe_fail(rc[(size_t)p], "partition " + std::to_string(p) + ": " + msg[(size_t)p]);
    return QE_OK;
}

extern "C" int qe_partition_rows(int64_t nrows, int nparts, int part, int64_t *begin, int64_t *end)
{
    if (nrows < 0 || nparts <= 0 || part < 0 || part >= nparts) return qe_fail(QE_ERR_INVALID_ARGUMENT, "bad partition arguments");
    int64_t per = (nrows + nparts - 1) / nparts;
    per = (per + QE_AGG_SUPER_ROWS - 1) / QE_AGG_SUPER_ROWS * QE_AGG_SUPER_ROWS;       // sums do not depend on nparts
    const int64_t b = std::min<int64_t>(nrows, (int64_t)part * per);
    if (begin) *begin = b;
    if (end) *end = std::min<int64_t>(nrows, b + per);
    return QE_OK;
}

extern "C" int qe_mgpu_create(const int *devices, int ndevices, qe_mgpu **out)
{
    if (!out) return qe_fail(QE_ERR_INVALID_ARGUMENT, "out is NULL");
    *out = nullptr;
    if (!devices || ndevices <= 0 || ndevices > 64) return qe_fail(QE_ERR_INVALID_ARGUMENT, "bad device list");
    std::unique_ptr<qe_mgpu> m(new qe_mgpu());
    for (int i = 0; i < ndevices; i++) {
        qe_ctx *c = nullptr;
        const int rc = qe_ctx_create(devices[i], &c);
        if (rc) {
            const std::string why = qe_last_error();
            for (qe_ctx *x : m->ctx) qe_ctx_destroy(x);
            return qe_fail(rc, "device " + std::to_string(devices[i]) + ": " + why);
        }
        m->devices.push_back(devices[i]);
        m->ctx.push_back(c);
        m->workers.emplace_back(new Worker());
    }
    *out = m.release();
    return QE_OK;
}

extern "C" int qe_mgpu_destroy(qe_mgpu *m)
{
    if (!m) return QE_OK;
    m->workers.clear();                          // joins the threads
    for (qe_ctx *c : m->ctx) qe_ctx_destroy(c);
    delete m;
    return QE_OK;
}

extern "C" int qe_mgpu_num_devices(const qe_mgpu *m) { return m ? (int)m->ctx.size() : 0; }

extern "C" qe_ctx *qe_mgpu_ctx(qe_mgpu *m, int part)
{
    return m && part >= 0 && part < (int)m->ctx.size() ? m->ctx[(size_t)part] : nullptr;
}

extern "C" int qe_mgpu_set_option(qe_mgpu *m, const char *key, int64_t value)
{
    if (!m) return qe_fail(QE_ERR_INVALID_ARGUMENT, "group is NULL");
    for (qe_ctx *c : m->ctx) {
        const int rc = qe_ctx_set_option(c, key, value);
        if (rc) return rc;
    }
    return QE_OK;
}

// ------------------------------------------------------------------- table

extern "C" int qe_mtable_create(qe_mgpu *m, int ncols, const qe_col_desc *cols, int64_t nrows, qe_mtable **out)
{
    if (!out) return qe_fail(QE_ERR_INVALID_ARGUMENT, "out is NULL");
    *out = nullptr;
    if (!m || !cols || ncols <= 0 || nrows < 0) return qe_fail(QE_ERR_INVALID_ARGUMENT, "bad table arguments");
    std::unique_ptr<qe_mtable> t(new qe_mtable());
    t->m = m;
    t->nrows = nrows;
    t->cols.assign(cols, cols + ncols);
    const int n = (int)m->ctx.size();
    t->parts.assign((size_t)n, nullptr);
    t->begin.assign((size_t)n, 0);
    t->end.assign((size_t)n, 0);
    for (int p = 0; p < n; p++) qe_partition_rows(nrows, n, p, &t->begin[(size_t)p], &t->end[(size_t)p]);
    qe_mtable *tp = t.get();
    const int rc = fan_out(m, [=](int p) {
        return qe_table_create(m->ctx[(size_t)p], ncols, tp->cols.data(), tp->end[(size_t)p] - tp->begin[(size_t)p], &tp->parts[(size_t)p]);
    });
    if (rc) {
        const std::string why = qe_last_error();
        for (qe_table *x : t->parts) qe_table_destroy(x);
        return qe_fail(rc, why);
    }
    *out = t.release();
    return QE_OK;
}

extern "C" int qe_mtable_destroy(qe_mtable *t)
{
    if (!t) return QE_OK;
    for (qe_table *x : t->parts) qe_table_destroy(x);
    delete t;
    return QE_OK;
}

extern "C" qe_table *qe_mtable_part(qe_mtable *t, int part)
{
    return t && part >= 0 && part < (int)t->parts.size() ? t->parts[(size_t)part] : nullptr;
}

extern "C" int qe_mtable_upload(qe_mtable *t, int col, const void *values, const uint8_t *validity)
{
    if (!t) return qe_fail(QE_ERR_INVALID_ARGUMENT, "table is NULL");
    if (col < 0 || col >= (int)t->cols.size()) return qe_fail(QE_ERR_INVALID_ARGUMENT, "column index out of range");
    const int type = t->cols[(size_t)col].type;
    const size_t w = width_of(type);
    return fan_out(t->m, [=](int p) {
        const int64_t b = t->begin[(size_t)p], n = t->end[(size_t)p] - b;      // b is a multiple of 2^20: whole bitmap bytes
        const void *v = !values ? nullptr : type == QE_TYPE_BOOLEAN ? (const void *)((const char *)values + b / 8)
                                                                      : (const void *)((const char *)values + (size_t)b * w);
        int rc = qe_table_upload(t->parts[(size_t)p], col, v, validity ? validity + b / 8 : nullptr, 0, n);
        if (!rc) rc = qe_ctx_synchronize(t->m->ctx[(size_t)p]);                  // the caller's buffers are free again on return
        return rc;
    });
}

extern "C" int qe_mtable_generate(qe_mtable *t, int col, int kind, uint64_t seed, double p0, double p1, double null_rate)
{
    if (!t) return qe_fail(QE_ERR_INVALID_ARGUMENT, "table is NULL");
    return fan_out(t->m, [=](int p) {
        return qe_table_generate(t->parts[(size_t)p], col, kind, seed, p0, p1, null_rate, t->begin[(size_t)p]);
    });
}

// ------------------------------------------------------------------ running

static int prepare_result(qe_mgpu *m, const qe_mtable *t, qe_mresult **result, qe_mresult **out)
{
    if (!m || !t || !result) return qe_fail(QE_ERR_INVALID_ARGUMENT, "NULL argument");
    if (t->m != m) return qe_fail(QE_ERR_INVALID_ARGUMENT, "table belongs to another device group");
    qe_mresult *r = *result;
    if (r && r->m != m) return qe_fail(QE_ERR_INVALID_ARGUMENT, "result belongs to another device group");
    if (!r) {
        r = new qe_mresult();
        r->m = m;
        r->parts.assign(m->ctx.size(), nullptr);
    }
    r->begin = t->begin;
    r->is_agg = r->is_group = false;
    r->agg_fns.clear();
    r->g_rows = 0;
    *out = r;
    return QE_OK;
}

extern "C" int qe_mgpu_pipeline_run(qe_mgpu *m, const qe_mtable *t, const qe_pipeline_desc *desc, qe_mresult **result)
{
    if (!desc) return qe_fail(QE_ERR_INVALID_ARGUMENT, "desc is NULL");
    if (desc->row_begin != 0 || desc->row_end >= 0) return qe_fail(QE_ERR_INVALID_ARGUMENT, "a sharded run covers the whole table: row_begin 0, row_end -1");
    qe_mresult *r = nullptr;
    int rc = prepare_result(m, t, result, &r);
    if (rc) return rc;
    r->is_agg = desc->nagg > 0;
    if (r->is_agg) r->agg_fns.assign(desc->agg_fns, desc->agg_fns + desc->nagg);
    const qe_pipeline_desc d = *desc;
    rc = fan_out(m, [=](int p) { return qe_pipeline_run_ex(m->ctx[(size_t)p], t->parts[(size_t)p], &d, &r->parts[(size_t)p]); });
    if (rc) {
        if (*result != r) { const std::string why = qe_last_error(); qe_mresult_destroy(r); return qe_fail(rc, why); }
        return rc;
    }
    *result = r;
    return QE_OK;
}

extern "C" int qe_mgpu_pipeline_stream(qe_mgpu *m, qe_mtable *t, const qe_stream_desc *descs)
{
    if (!m || !t || !descs) return qe_fail(QE_ERR_INVALID_ARGUMENT, "NULL argument");
    if (t->m != m) return qe_fail(QE_ERR_INVALID_ARGUMENT, "table belongs to another device group");
    return fan_out(m, [=](int p) { return qe_pipeline_stream(m->ctx[(size_t)p], t->parts[(size_t)p], &descs[p]); });
}

extern "C" int qe_mresult_destroy(qe_mresult *r)
{
    if (!r) return QE_OK;
    for (qe_result *x : r->parts) qe_result_destroy(x);
    delete r;
    return QE_OK;
}

extern "C" qe_result *qe_mresult_part(qe_mresult *r, int part)
{
    return r && part >= 0 && part < (int)r->parts.size() ? r->parts[(size_t)part] : nullptr;
}

extern "C" int qe_mresult_part_rows(qe_mresult *r, int part, int64_t *nrows)
{
    if (!r || part < 0 || part >= (int)r->parts.size()) return qe_fail(QE_ERR_INVALID_ARGUMENT, "partition index out of range");
    return qe_result_num_rows(r->parts[(size_t)part], nrows);
}

extern "C" int qe_mresult_num_rows(qe_mresult *r, int64_t *nrows)
{
    if (!r) return qe_fail(QE_ERR_INVALID_ARGUMENT, "result is NULL");
    int64_t total = 0;
    if (r->is_group) total = r->g_rows;
    else if (r->is_agg) {
        for (qe_result *x : r->parts) { int64_t n = 0; const int rc = qe_result_num_rows(x, &n); if (rc) return rc; }
        total = 1;
    } else {
        for (qe_result *x : r->parts) {
            int64_t n = 0;
            const int rc = qe_result_num_rows(x, &n);
            if (rc) return rc;
            total += n;
        }
    }
    if (nrows) *nrows = total;
    return QE_OK;
}

extern "C" int qe_mresult_num_columns(const qe_mresult *r, int *ncols)
{
    if (!r || !ncols) return qe_fail(QE_ERR_INVALID_ARGUMENT, "NULL argument");
    if (r->is_group) { *ncols = (int)r->g_types.size(); return QE_OK; }
    return qe_result_num_columns(r->parts[0], ncols);
}

extern "C" int qe_mresult_agg(qe_mresult *r, int k, double *value, int *is_null)
{
    if (!r) return qe_fail(QE_ERR_INVALID_ARGUMENT, "result is NULL");
    if (!r->is_agg) return qe_fail(QE_ERR_STATE, "this result is not a global aggregation");
    if (k < 0 || k >= (int)r->agg_fns.size()) return qe_fail(QE_ERR_INVALID_ARGUMENT, "aggregate index out of range");
    // the super-tile partials of all partitions, in row order, folded one by one: the same additions in the same
    // order whatever the number of devices
    std::vector<qe_agg_partial> list;
    for (qe_result *x : r->parts) {
        int64_t n = 0;
        int rc = qe_result_agg_partials(x, k, nullptr, 0, &n);
        if (rc) return rc;
        const size_t at = list.size();
        list.resize(at + (size_t)n);
        if (n && (rc = qe_result_agg_partials(x, k, list.data() + at, n, &n))) return rc;
    }
    return qe_agg_combine(r->agg_fns[(size_t)k], list.data(), (int)list.size(), value, is_null);
}

extern "C" int qe_mresult_column(qe_mresult *r, int k, void *dst_values, uint8_t *dst_validity)
{
    if (!r) return qe_fail(QE_ERR_INVALID_ARGUMENT, "result is NULL");
    if (r->is_group) {
        if (k < 0 || k >= (int)r->g_types.size()) return qe_fail(QE_ERR_INVALID_ARGUMENT, "column index out of range");
        if (dst_values) memcpy(dst_values, r->g_values[(size_t)k].data(), r->g_values[(size_t)k].size());
        if (dst_validity) memcpy(dst_validity, r->g_valid[(size_t)k].data(), r->g_valid[(size_t)k].size());
        return QE_OK;
    }
    if (r->is_agg) {
        double v = 0.0;
        int is_null = 0;
        const int rc = qe_mresult_agg(r, k, &v, &is_null);
        if (rc) return rc;
        if (dst_values) *(double *)dst_values = v;
        if (dst_validity) *dst_validity = is_null ? 0 : 1;
        return QE_OK;
    }
    const int n = (int)r->parts.size();
    std::vector<int64_t> rows((size_t)n, 0), off((size_t)n + 1, 0);
    for (int p = 0; p < n; p++) {
        const int rc = qe_result_num_rows(r->parts[(size_t)p], &rows[(size_t)p]);
        if (rc) return rc;
        off[(size_t)p + 1] = off[(size_t)p] + rows[(size_t)p];
    }
    int type = 0, has_validity = 0;
    int rc = qe_result_column_type(r->parts[0], k, &type, &has_validity);
    if (rc) return rc;
    const size_t w = width_of(type);
    const bool bits_v = type == QE_TYPE_BOOLEAN && dst_values, bits_n = dst_validity != nullptr;
    std::vector<std::vector<uint8_t>> tv((size_t)n), tn((size_t)n);
    // every device copies its rows at the same time; byte-addressed values land in place, bit columns are spliced below
    rc = fan_out(r->m, [&](int p) {
        if (rows[(size_t)p] == 0) return (int)QE_OK;
        if (bits_v) tv[(size_t)p].assign((size_t)(rows[(size_t)p] + 7) / 8 + 1, 0);
        if (bits_n) tn[(size_t)p].assign((size_t)(rows[(size_t)p] + 7) / 8 + 1, 0);
        void *v = !dst_values ? nullptr : bits_v ? (void *)tv[(size_t)p].data() : (void *)((char *)dst_values + (size_t)off[(size_t)p] * w);
        return qe_result_column(r->parts[(size_t)p], k, v, bits_n ? tn[(size_t)p].data() : nullptr);
    });
    if (rc) return rc;
    const int64_t total = off[(size_t)n];
    if (bits_v) memset(dst_values, 0, (size_t)(total + 7) / 8);
    if (bits_n) memset(dst_validity, 0, (size_t)(total + 7) / 8);
    for (int p = 0; p < n; p++) {
        if (bits_v && rows[(size_t)p]) append_bits((uint8_t *)dst_values, off[(size_t)p], tv[(size_t)p].data(), rows[(size_t)p]);
        if (bits_n && rows[(size_t)p]) append_bits(dst_validity, off[(size_t)p], tn[(size_t)p].data(), rows[(size_t)p]);
    }
    return QE_OK;
}

extern "C" int qe_mresult_selection(qe_mresult *r, uint64_t *dst)
{
    if (!r || !dst) return qe_fail(QE_ERR_INVALID_ARGUMENT, "NULL argument");
    if (r->is_agg || r->is_group) return qe_fail(QE_ERR_STATE, "this result has no selection vector");
    const int n = (int)r->parts.size();
    std::vector<int64_t> rows((size_t)n, 0), off((size_t)n + 1, 0);
    for (int p = 0; p < n; p++) {
        const int rc = qe_result_num_rows(r->parts[(size_t)p], &rows[(size_t)p]);
        if (rc) return rc;
        off[(size_t)p + 1] = off[(size_t)p] + rows[(size_t)p];
    }
    return fan_out(r->m, [&](int p) {
        const int64_t np = rows[(size_t)p];
        if (np == 0) return (int)QE_OK;
        // partition-relative 32-bit indices land in the upper half of the partition's 64-bit slots and are widened in place
        uint64_t *out = dst + off[(size_t)p];
        uint32_t *tmp = (uint32_t *)out + np;
        const int rc = qe_result_selection(r->parts[(size_t)p], tmp);
        if (rc) return rc;
        const uint64_t base = (uint64_t)r->begin[(size_t)p];
        for (int64_t i = 0; i < np; i++) out[i] = base + tmp[i];
        return (int)QE_OK;
    });
}

// ------------------------------------------------------------------ group-by
// Every device groups its own rows (qe_groupby_run); the per-partition groups -- a handful of fixed-width records --
// are merged by key here: partials of one group folded in partition order (qe_agg_combine), groups ordered by the
// smallest global row at which they appear, which is the order a single scan would have produced (LinkedHashMap,
// M/operator/GroupByAggregationOperator.kt:22,29-36).

extern "C" int qe_groupby_merge(const qe_group_part *parts, int nparts, int nkeys, const int32_t *agg_fns, int nagg,
                                int64_t capacity, uint64_t *out_keys, uint8_t *out_key_null, int64_t *out_first_row,
                                double *out_values, uint8_t *out_null, int64_t *out_ngroups)
{
    if (!parts || nparts < 0 || nkeys < 0 || nkeys > 8 || nagg < 0 || (nagg && !agg_fns) || !out_ngroups)
        return qe_fail(QE_ERR_INVALID_ARGUMENT, "bad merge arguments");
    const size_t krec = (size_t)nkeys * 9;                    // per key 8 value bytes + 1 null byte
    struct Group { int64_t first; std::vector<qe_agg_partial> parts; };
    std::unordered_map<std::string, size_t> index;
    std::vector<std::string> keys;
    std::vector<Group> groups;
    for (int p = 0; p < nparts; p++) {
        const qe_group_part &gp = parts[p];
        if (gp.ngroups < 0 || (gp.ngroups > 0 && ((nkeys && (!gp.keys || !gp.key_null)) || !gp.first_row || (nagg && !gp.partials))))
            return qe_fail(QE_ERR_INVALID_ARGUMENT, "bad group partition");
        for (int64_t g = 0; g < gp.ngroups; g++) {
            std::string key(krec, '\0');
            for (int k = 0; k < nkeys; k++) {
                const bool null = gp.key_null[(size_t)g * (size_t)nkeys + (size_t)k] != 0;
                const uint64_t bits = null ? 0 : gp.keys[(size_t)g * (size_t)nkeys + (size_t)k];
                memcpy(&key[(size_t)k * 9], &bits, 8);
                key[(size_t)k * 9 + 8] = null ? 1 : 0;
            }
            auto it = index.find(key);
            size_t gi;
            if (it == index.end()) {
                gi = groups.size();
                index.emplace(key, gi);
                keys.push_back(key);
                groups.push_back(Group{gp.first_row[g], {}});
            } else {
                gi = it->second;
                groups[gi].first = std::min(groups[gi].first, gp.first_row[g]);
            }
            for (int j = 0; j < nagg; j++) groups[gi].parts.push_back(gp.partials[(size_t)g * (size_t)nagg + (size_t)j]);   // partition order
        }
    }
    const int64_t ng = (int64_t)groups.size();
    *out_ngroups = ng;
    if (ng > capacity) return qe_fail(QE_ERR_CAPACITY, "merged group-by has " + std::to_string(ng) + " groups, the output arrays hold " + std::to_string(capacity));
    std::vector<size_t> order((size_t)ng);
    for (size_t i = 0; i < order.size(); i++) order[i] = i;
    std::sort(order.begin(), order.end(), [&](size_t a, size_t b) { return groups[a].first < groups[b].first; });
    std::vector<qe_agg_partial> one;
    for (int64_t o = 0; o < ng; o++) {
        const size_t gi = order[(size_t)o];
        for (int k = 0; k < nkeys; k++) {
            if (out_keys) memcpy(&out_keys[(size_t)o * (size_t)nkeys + (size_t)k], &keys[gi][(size_t)k * 9], 8);
            if (out_key_null) out_key_null[(size_t)o * (size_t)nkeys + (size_t)k] = (uint8_t)keys[gi][(size_t)k * 9 + 8];
        }
        if (out_first_row) out_first_row[o] = groups[gi].first;
        const std::vector<qe_agg_partial> &ps = groups[gi].parts;
        const int nseen = nagg ? (int)(ps.size() / (size_t)nagg) : 0;
        one.resize((size_t)nseen);
        for (int j = 0; j < nagg; j++) {
            for (int s = 0; s < nseen; s++) one[(size_t)s] = ps[(size_t)s * (size_t)nagg + (size_t)j];
            double v = 0.0;
            int is_null = 0;
            const int rc = qe_agg_combine(agg_fns[j], one.data(), nseen, &v, &is_null);
            if (rc) return rc;
            if (out_values) out_values[(size_t)o * (size_t)nagg + (size_t)j] = is_null ? 0.0 : v;
            if (out_null) out_null[(size_t)o * (size_t)nagg + (size_t)j] = is_null ? 1 : 0;
        }
    }
    return QE_OK;
}

extern "C" int qe_mgpu_groupby_run(qe_mgpu *m, const qe_mtable *t, const qe_groupby_desc *desc, qe_mresult **result)
{
    if (!desc) return qe_fail(QE_ERR_INVALID_ARGUMENT, "desc is NULL");
    if (desc->row_begin != 0 || desc->row_end >= 0) return qe_fail(QE_ERR_INVALID_ARGUMENT, "a sharded run covers the whole table: row_begin 0, row_end -1");
    qe_mresult *r = nullptr;
    int rc = prepare_result(m, t, result, &r);
    if (rc) return rc;
    struct Drop { qe_mresult *r; qe_mresult **slot; ~Drop() { if (*slot != r) qe_mresult_destroy(r); } } drop{r, result};
    const qe_groupby_desc d = *desc;
    rc = fan_out(m, [=](int p) { return qe_groupby_run(m->ctx[(size_t)p], t->parts[(size_t)p], &d, &r->parts[(size_t)p]); });
    if (rc) return rc;

    const int nparts = (int)r->parts.size(), nkeys = desc->nkeys, nagg = desc->nagg, ncols = nkeys + nagg;
    r->is_group = true;
    r->g_nkeys = nkeys;
    r->g_types.assign((size_t)ncols, QE_TYPE_DOUBLE);
    for (int k = 0; k < nkeys; k++) {
        int type = 0, hv = 0;
        if ((rc = qe_result_column_type(r->parts[0], k, &type, &hv))) return rc;
        r->g_types[(size_t)k] = type;
    }
    // every partition's groups as fixed-width records: key bits, null flags, global first row, partials
    struct PartData { std::vector<uint64_t> keys; std::vector<uint8_t> nulls; std::vector<int64_t> first; std::vector<qe_agg_partial> partials; };
    std::vector<PartData> data((size_t)nparts);
    std::vector<qe_group_part> gparts((size_t)nparts);
    int64_t total = 0;
    for (int p = 0; p < nparts; p++) {
        qe_result *x = r->parts[(size_t)p];
        PartData &pd = data[(size_t)p];
        int64_t ng = 0;
        if ((rc = qe_result_num_rows(x, &ng))) return rc;
        total += ng;
        pd.keys.assign((size_t)ng * (size_t)std::max(nkeys, 1), 0);
        pd.nulls.assign((size_t)ng * (size_t)std::max(nkeys, 1), 0);
        pd.first.assign((size_t)ng, 0);
        pd.partials.assign((size_t)ng * (size_t)std::max(nagg, 1), qe_agg_partial());
        if (ng > 0) {
            std::vector<uint8_t> kv((size_t)ng * 8 + 8), kn((size_t)(ng + 7) / 8 + 1);
            for (int k = 0; k < nkeys; k++) {
                std::fill(kv.begin(), kv.end(), 0);
                std::fill(kn.begin(), kn.end(), 0);
                if ((rc = qe_result_column(x, k, kv.data(), kn.data()))) return rc;
                const int type = r->g_types[(size_t)k];
                for (int64_t g = 0; g < ng; g++) {
                    const bool valid = (kn[(size_t)(g / 8)] >> (g % 8)) & 1u;
                    uint64_t bits = 0;
                    if (valid) {
                        if (type == QE_TYPE_BOOLEAN) bits = (kv[(size_t)(g / 8)] >> (g % 8)) & 1u;
                        else if (width_of(type) == 8) memcpy(&bits, kv.data() + (size_t)g * 8, 8);
                        else { uint32_t v32; memcpy(&v32, kv.data() + (size_t)g * 4, 4); bits = v32; }
                    }
                    pd.keys[(size_t)g * (size_t)nkeys + (size_t)k] = bits;
                    pd.nulls[(size_t)g * (size_t)nkeys + (size_t)k] = valid ? 0 : 1;
                }
            }
            if ((rc = qe_result_group_first_rows(x, pd.first.data()))) return rc;
            for (int64_t g = 0; g < ng; g++) {
                pd.first[(size_t)g] += r->begin[(size_t)p];
                for (int j = 0; j < nagg; j++)
                    if ((rc = qe_result_group_partial(x, g, j, &pd.partials[(size_t)g * (size_t)nagg + (size_t)j]))) return rc;
            }
        }
        gparts[(size_t)p] = qe_group_part{ng, pd.keys.data(), pd.nulls.data(), pd.first.data(), pd.partials.data()};
    }
    std::vector<uint64_t> mk((size_t)total * (size_t)std::max(nkeys, 1) + 1);
    std::vector<uint8_t> mn((size_t)total * (size_t)std::max(nkeys, 1) + 1), vn((size_t)total * (size_t)std::max(nagg, 1) + 1);
    std::vector<double> mv((size_t)total * (size_t)std::max(nagg, 1) + 1);
    int64_t ng = 0;
    if ((rc = qe_groupby_merge(gparts.data(), nparts, nkeys, desc->agg_fns, nagg, total, mk.data(), mn.data(), nullptr, mv.data(), vn.data(), &ng)))
        return rc;
    r->g_rows = ng;
    r->g_values.assign((size_t)ncols, std::vector<uint8_t>());
    r->g_valid.assign((size_t)ncols, std::vector<uint8_t>());
    const size_t vbytes = (size_t)(ng + 7) / 8;
    for (int c = 0; c < ncols; c++) {
        const int type = r->g_types[(size_t)c];
        r->g_values[(size_t)c].assign(type == QE_TYPE_BOOLEAN ? vbytes : (size_t)ng * width_of(type), 0);
        r->g_valid[(size_t)c].assign(vbytes, 0);
    }
    for (int64_t o = 0; o < ng; o++) {
        for (int k = 0; k < nkeys; k++) {
            if (mn[(size_t)o * (size_t)nkeys + (size_t)k]) continue;
            r->g_valid[(size_t)k][(size_t)(o / 8)] |= (uint8_t)(1u << (o % 8));
            const uint64_t bits = mk[(size_t)o * (size_t)nkeys + (size_t)k];
            const int type = r->g_types[(size_t)k];
            uint8_t *dst = r->g_values[(size_t)k].data();
            if (type == QE_TYPE_BOOLEAN) { if (bits) dst[o / 8] |= (uint8_t)(1u << (o % 8)); }
            else if (width_of(type) == 8) memcpy(dst + (size_t)o * 8, &bits, 8);
            else { const uint32_t v32 = (uint32_t)bits; memcpy(dst + (size_t)o * 4, &v32, 4); }
        }
        for (int j = 0; j < nagg; j++) {
            if (vn[(size_t)o * (size_t)nagg + (size_t)j]) continue;
            memcpy(r->g_values[(size_t)(nkeys + j)].data() + (size_t)o * 8, &mv[(size_t)o * (size_t)nagg + (size_t)j], 8);
            r->g_valid[(size_t)(nkeys + j)][(size_t)(o / 8)] |= (uint8_t)(1u << (o % 8));
        }
    }
    *result = r;
    return QE_OK;
}
