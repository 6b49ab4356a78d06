// qe_jit.h -- run-time compilation of generated pipeline programs with NVRTC
// (loaded lazily with dlopen, so libqegpu.so itself has no link-time
// dependency on it) and loading of the resulting sm_100a cubin through the
// CUDA runtime's library API.
#ifndef QE_JIT_H
#define QE_JIT_H

#include <cuda_runtime.h>

#include <string>
#include <vector>

namespace qe {

// Text of the device headers, embedded at build time (qe_embedded_src.cpp is
// generated from qe_device.cuh / qe_ops.h by the build script).
extern const char *const kEmbeddedDeviceHeader;
extern const char *const kEmbeddedOpsHeader;

struct JitKernel {
    cudaLibrary_t library = nullptr;
    cudaKernel_t kernel = nullptr;
    int blocks_per_sm = 1;
    int regs = 0;
    size_t local_bytes = 0; // per-thread local memory (register spills)
    size_t dyn_smem = 0;    // dynamic shared memory per CTA
};

// True when libnvrtc could be loaded.  `why` receives the reason otherwise.
bool jit_available(std::string *why);

// 256-bit global loads/stores (VEC = 4 tile shapes) need NVRTC >= 12.9.
bool jit_wide_loads_supported();
std::string jit_version();

// Compile `source` (which includes "qe_device.cuh") to a cubin for sm_100a.
// Does not need a GPU.  Returns false and fills `log` on failure.
bool jit_compile(const std::string &source, std::vector<char> &cubin, std::string &log,
                 const std::vector<std::string> &extra_options = std::vector<std::string>());

// Load a cubin on the current device and look up the entry point.
bool jit_load(const std::vector<char> &cubin, const char *entry, size_t dyn_smem, int block_threads, JitKernel &out, std::string &err);
void jit_unload(JitKernel &k);

}  // namespace qe
#endif
