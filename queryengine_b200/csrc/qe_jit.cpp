// qe_jit.cpp -- see qe_jit.h
#include "qe_jit.h"

#include <dlfcn.h>

#include <cstdlib>
#include <mutex>

namespace qe {
namespace {

// The handful of NVRTC entry points used, resolved with dlsym.
typedef struct _nvrtcProgram *nvrtcProgram;
typedef int nvrtcResult;

struct Nvrtc {
    void *handle = nullptr;
    std::string why;
    nvrtcResult (*CreateProgram)(nvrtcProgram *, const char *, const char *, int, const char *const *, const char *const *) = nullptr;
    nvrtcResult (*CompileProgram)(nvrtcProgram, int, const char *const *) = nullptr;
    nvrtcResult (*DestroyProgram)(nvrtcProgram *) = nullptr;
    nvrtcResult (*GetCUBINSize)(nvrtcProgram, size_t *) = nullptr;
    nvrtcResult (*GetCUBIN)(nvrtcProgram, char *) = nullptr;
    nvrtcResult (*GetProgramLogSize)(nvrtcProgram, size_t *) = nullptr;
    nvrtcResult (*GetProgramLog)(nvrtcProgram, char *) = nullptr;
    const char *(*GetErrorString)(nvrtcResult) = nullptr;
    nvrtcResult (*Version)(int *, int *) = nullptr;
    int major = 0, minor = 0;
};

Nvrtc &nvrtc()
{
    static Nvrtc n;
    static std::once_flag once;
    std::call_once(once, [] {
        // tests / deployments without a compiler: behave as if libnvrtc were missing (AOT and disk cubins still load)
        if (const char *off = getenv("QE_DISABLE_NVRTC"); off && *off && *off != '0') { n.why = "NVRTC disabled by QE_DISABLE_NVRTC"; return; }
        // Absolute paths of the toolkit first: a bare soname would resolve to whatever libnvrtc the
        // process already has loaded (PyTorch bundles an older one whose ptxas lacks the 256-bit
        // loads of sm_100), and loading by path gets the toolkit's own copy beside it.
        const char *candidates[] = {
            getenv("QE_NVRTC_PATH"),
            "/usr/local/cuda/lib64/libnvrtc.so.12",
            "/usr/local/cuda/targets/x86_64-linux/lib/libnvrtc.so.12",
            "/usr/local/cuda-12.9/targets/x86_64-linux/lib/libnvrtc.so.12",
            "libnvrtc.so.12",
            "libnvrtc.so",
            "libnvrtc.so.13",
        };
        for (const char *c : candidates) {
            if (!c || !*c) continue;
            n.handle = dlopen(c, RTLD_NOW | RTLD_LOCAL);
            if (n.handle) break;
        }
        if (!n.handle) { n.why = "libnvrtc.so.12 not found (set QE_NVRTC_PATH)"; return; }
#define QE_SYM(name) \
        *(void **)(&n.name) = dlsym(n.handle, "nvrtc" #name); \
        if (!n.name) { n.why = "nvrtc" #name " missing from libnvrtc"; n.handle = nullptr; return; }
        QE_SYM(CreateProgram) QE_SYM(CompileProgram) QE_SYM(DestroyProgram) QE_SYM(GetCUBINSize) QE_SYM(GetCUBIN)
        QE_SYM(GetProgramLogSize) QE_SYM(GetProgramLog) QE_SYM(GetErrorString) QE_SYM(Version)
#undef QE_SYM
        n.Version(&n.major, &n.minor);
    });
    return n;
}

}  // namespace

bool jit_available(std::string *why)
{
    Nvrtc &n = nvrtc();
    if (!n.handle && why) *why = n.why;
    return n.handle != nullptr;
}

bool jit_wide_loads_supported()
{
    Nvrtc &n = nvrtc();
    // ld.global.v4.f64 / .v4.b64 (LDG.E.256) needs PTX ISA 8.8 = CUDA 12.9.  Without any NVRTC the shapes of the
    // ahead-of-time cubins apply (built with the toolkit's 12.9), so the answer is yes.
    return !n.handle || n.major > 12 || (n.major == 12 && n.minor >= 9);
}

std::string jit_version()
{
    Nvrtc &n = nvrtc();
    return n.handle ? std::to_string(n.major) + "." + std::to_string(n.minor) : std::string("unavailable");
}

bool jit_compile(const std::string &source, std::vector<char> &cubin, std::string &log,
                 const std::vector<std::string> &extra_options)
{
    Nvrtc &n = nvrtc();
    if (!n.handle) { log = n.why; return false; }
    const char *headers[] = {kEmbeddedDeviceHeader, kEmbeddedOpsHeader};
    const char *names[] = {"qe_device.cuh", "qe_ops.h"};
    nvrtcProgram prog = nullptr;
    nvrtcResult rc = n.CreateProgram(&prog, source.c_str(), "qe_pipeline.cu", 2, headers, names);
    if (rc != 0) { log = std::string("nvrtcCreateProgram: ") + n.GetErrorString(rc); return false; }
    // -fmad=false: a*b+c must stay a separate multiply and add, as on the JVM
    std::vector<const char *> opts = {"--gpu-architecture=sm_100a", "--std=c++17", "--fmad=false", "-lineinfo"};
    for (const std::string &o : extra_options) opts.push_back(o.c_str());
    rc = n.CompileProgram(prog, (int)opts.size(), opts.data());
    size_t logsz = 0;
    n.GetProgramLogSize(prog, &logsz);
    if (logsz > 1) {
        std::vector<char> buf(logsz);
        n.GetProgramLog(prog, buf.data());
        log.assign(buf.data());
    }
    if (rc != 0) {
        log = std::string("nvrtcCompileProgram: ") + n.GetErrorString(rc) + "\n" + log;
        n.DestroyProgram(&prog);
        return false;
    }
    size_t sz = 0;
    rc = n.GetCUBINSize(prog, &sz);
    if (rc != 0 || sz == 0) {
        log = "nvrtcGetCUBINSize failed";
        n.DestroyProgram(&prog);
        return false;
    }
    cubin.resize(sz);
    n.GetCUBIN(prog, cubin.data());
    n.DestroyProgram(&prog);
    return true;
}

bool jit_load(const std::vector<char> &cubin, const char *entry, size_t dyn_smem, int block_threads, JitKernel &out, std::string &err)
{
    cudaError_t e = cudaLibraryLoadData(&out.library, cubin.data(), nullptr, nullptr, 0, nullptr, nullptr, 0);
    if (e != cudaSuccess) { err = std::string("cudaLibraryLoadData: ") + cudaGetErrorString(e); return false; }
    e = cudaLibraryGetKernel(&out.kernel, out.library, entry);
    if (e != cudaSuccess) {
        err = std::string("cudaLibraryGetKernel: ") + cudaGetErrorString(e);
        cudaLibraryUnload(out.library);
        out.library = nullptr;
        return false;
    }
    out.dyn_smem = dyn_smem;
    if (dyn_smem > 0) {
        e = cudaFuncSetAttribute((const void *)out.kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn_smem);
        if (e != cudaSuccess) {
            err = std::string("cudaFuncSetAttribute(MaxDynamicSharedMemorySize): ") + cudaGetErrorString(e);
            cudaLibraryUnload(out.library);
            out.library = nullptr;
            return false;
        }
    }
    int n = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, (const void *)out.kernel, block_threads, dyn_smem) == cudaSuccess && n > 0)
        out.blocks_per_sm = n;
    else
        (void)cudaGetLastError();
    cudaFuncAttributes fa;
    if (cudaFuncGetAttributes(&fa, (const void *)out.kernel) == cudaSuccess) { out.regs = fa.numRegs; out.local_bytes = fa.localSizeBytes; }
    else (void)cudaGetLastError();
    return true;
}

void jit_unload(JitKernel &k)
{
    if (k.library) cudaLibraryUnload(k.library);
    k.library = nullptr;
    k.kernel = nullptr;
}

}  // namespace qe
