// qe_cubin_cache.cpp -- see qe_cubin_cache.h
#include "qe_cubin_cache.h"

#include <sys/stat.h>
#include <sys/types.h>
#include <unistd.h>

#include <cerrno>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <unordered_map>

#include "qe_jit.h"

namespace qe {
namespace {

// Two independent 64-bit FNV-1a style lanes; the key only has to separate a few thousand kernels, the file also
// stores the key of what it holds, and a cubin that does not contain the entry point fails to load cleanly.
struct Hash128 {
    uint64_t a = 0xcbf29ce484222325ULL, b = 0x84222325cbf29ce4ULL;
    void feed(const void *p, size_t n)
    {
        const unsigned char *s = (const unsigned char *)p;
        for (size_t i = 0; i < n; i++) {
            a = (a ^ s[i]) * 0x100000001b3ULL;
            b = (b ^ (s[i] + 0x9eULL)) * 0x9e3779b97f4a7c15ULL;
            b ^= b >> 29;
        }
    }
    void feed(const std::string &s) { feed(s.data(), s.size()); const unsigned char z = 0; feed(&z, 1); }
    std::string hex() const
    {
        char buf[40];
        snprintf(buf, sizeof buf, "%016llx%016llx", (unsigned long long)a, (unsigned long long)b);
        return buf;
    }
};

struct Store {
    std::mutex mu;
    std::mutex compile_mu;            // one NVRTC compilation at a time: the second thread finds the first one's result
    std::unordered_map<std::string, CubinRef> mem;
    CubinStats stats;
    bool dir_known = false;
    std::string dir;
};

Store &store()
{
    static Store s;
    return s;
}

const Hash128 &header_hash()
{
    static const Hash128 h = [] {
        Hash128 x;
        x.feed(kEmbeddedDeviceHeader, strlen(kEmbeddedDeviceHeader));
        x.feed(kEmbeddedOpsHeader, strlen(kEmbeddedOpsHeader));
        x.feed(std::string("sm_100a|fmad=false|lineinfo|v2"));
        return x;
    }();
    return h;
}

bool make_dirs(const std::string &path)
{
    std::string cur;
    for (size_t i = 0; i <= path.size(); i++) {
        if (i == path.size() || path[i] == '/') {
            if (!cur.empty() && mkdir(cur.c_str(), 0755) != 0 && errno != EEXIST) return false;
        }
        if (i < path.size()) cur += path[i];
    }
    return true;
}

std::string resolve_dir()
{
    const char *e = getenv("QE_CACHE_DIR");
    std::string d;
    if (e) {
        d = e;
        if (d.empty() || d == "off" || d == "0") return "";
    } else if (const char *x = getenv("XDG_CACHE_HOME"); x && *x) {
        d = std::string(x) + "/qegpu";
    } else if (const char *h = getenv("HOME"); h && *h) {
        d = std::string(h) + "/.cache/qegpu";
    } else {
        return "";
    }
    return make_dirs(d) ? d : "";
}

const char kMagic[8] = {'Q', 'E', 'C', 'U', 'B', 'I', 'N', '2'};

bool read_file(const std::string &path, const std::string &key, std::vector<char> &cubin)
{
    FILE *f = fopen(path.c_str(), "rb");
    if (!f) return false;
    char magic[8], k[32];
    uint64_t size = 0;
    bool ok = fread(magic, 1, 8, f) == 8 && memcmp(magic, kMagic, 8) == 0 && fread(k, 1, 32, f) == 32 &&
              memcmp(k, key.data(), 32) == 0 && fread(&size, 8, 1, f) == 1 && size > 0 && size < (256u << 20);
    if (ok) {
        cubin.resize((size_t)size);
        ok = fread(cubin.data(), 1, (size_t)size, f) == (size_t)size;
    }
    fclose(f);
    return ok;
}

void write_file(const std::string &dir, const std::string &key, const std::vector<char> &cubin)
{
    char tmp[64];
    snprintf(tmp, sizeof tmp, "/.tmp.%d.%s", (int)getpid(), key.substr(0, 8).c_str());
    const std::string tpath = dir + tmp, path = dir + "/" + key + ".cubin";
    FILE *f = fopen(tpath.c_str(), "wb");
    if (!f) return;
    const uint64_t size = cubin.size();
    const bool ok = fwrite(kMagic, 1, 8, f) == 8 && fwrite(key.data(), 1, 32, f) == 32 && fwrite(&size, 8, 1, f) == 1 &&
                    fwrite(cubin.data(), 1, cubin.size(), f) == cubin.size();
    fclose(f);
    if (!ok || rename(tpath.c_str(), path.c_str()) != 0) unlink(tpath.c_str());   // readers see a whole file or none
}

}  // namespace

std::string cubin_key(const std::string &source, const std::vector<std::string> &options)
{
    Hash128 h = header_hash();
    h.feed(source);
    for (const std::string &o : options) h.feed(o);
    return h.hex();
}

std::string cubin_cache_dir()
{
    Store &s = store();
    std::lock_guard<std::mutex> g(s.mu);
    if (!s.dir_known) { s.dir = resolve_dir(); s.dir_known = true; }
    return s.dir;
}

CubinStats cubin_stats()
{
    Store &s = store();
    std::lock_guard<std::mutex> g(s.mu);
    return s.stats;
}

void cubin_forget_memory()
{
    Store &s = store();
    std::lock_guard<std::mutex> g(s.mu);
    s.mem.clear();
    s.dir_known = false;              // QE_CACHE_DIR is read again
}

int cubin_get(const std::string &source, const std::vector<std::string> &options, CubinRef &out, int *origin, std::string &log)
{
    Store &s = store();
    const std::string key = cubin_key(source, options);
    auto lookup_mem = [&]() -> bool {
        std::lock_guard<std::mutex> g(s.mu);
        auto it = s.mem.find(key);
        if (it == s.mem.end()) return false;
        out = it->second;
        s.stats.hits[CUBIN_MEMORY]++;
        if (origin) *origin = CUBIN_MEMORY;
        return true;
    };
    auto publish = [&](std::vector<char> &&bytes, int from) {
        CubinRef ref = std::make_shared<const std::vector<char>>(std::move(bytes));
        std::lock_guard<std::mutex> g(s.mu);
        s.mem[key] = ref;
        s.stats.hits[from]++;
        out = ref;
        if (origin) *origin = from;
    };
    if (lookup_mem()) return 0;
    if (!getenv("QE_DISABLE_AOT")) {
        for (size_t i = 0; i < kAotCount; i++) {
            if (key == kAotTable[i].hash) {
                publish(std::vector<char>((const char *)kAotTable[i].data, (const char *)kAotTable[i].data + kAotTable[i].size), CUBIN_AOT);
                return 0;
            }
        }
    }
    const std::string dir = cubin_cache_dir();
    if (!dir.empty()) {
        std::vector<char> bytes;
        if (read_file(dir + "/" + key + ".cubin", key, bytes)) {
            publish(std::move(bytes), CUBIN_DISK);
            return 0;
        }
    }
    std::string why;
    if (!jit_available(&why)) {
        log = why;
        return -1;
    }
    std::lock_guard<std::mutex> cg(s.compile_mu);
    if (lookup_mem()) return 0;       // another thread compiled it while this one waited
    std::vector<char> bytes;
    const auto t0 = std::chrono::steady_clock::now();
    if (!jit_compile(source, bytes, log, options)) return 6;
    const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    if (!dir.empty()) write_file(dir, key, bytes);
    publish(std::move(bytes), CUBIN_COMPILED);
    {
        std::lock_guard<std::mutex> g(s.mu);
        s.stats.compile_ms += ms;
    }
    return 0;
}

}  // namespace qe
