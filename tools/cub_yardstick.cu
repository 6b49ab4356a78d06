// cub_yardstick.cu -- same-box yardsticks for the filter/scan kernels (development tool, not part of libqegpu.so):
//   * pure-read and copy bandwidth of simple grid-stride kernels (what HBM gives a streaming kernel on this box)
//   * cub::DeviceSelect::If on the benchmark column (int64, `x < k`), compacted values only
//   * cub::DeviceSelect::If over (row index, value) pairs = the same outputs as qe_filter (selection vector + values)
// Build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o tools/cub_yardstick tools/cub_yardstick.cu
// Run:   tools/cub_yardstick [rows]
#include <cub/cub.cuh>
#include <cuda_runtime.h>
#include <thrust/iterator/counting_iterator.h>
#include <thrust/iterator/zip_iterator.h>
#include <thrust/tuple.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

__device__ __forceinline__ uint64_t splitmix64(uint64_t x)
{
    x += 0x9e3779b97f4a7c15ull;
    x = (x ^ (x >> 30)) * 0xbf58476d1ce4e5b9ull;
    x = (x ^ (x >> 27)) * 0x94d049bb133111ebull;
    return x ^ (x >> 31);
}

__global__ void gen(int64_t *v, int64_t n)
{
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        v[i] = (int64_t)(splitmix64(42ull ^ (uint64_t)i) >> 24);
}

__global__ void __launch_bounds__(256) read_kernel(const longlong4 *v, int64_t n4, unsigned long long *out)
{
    unsigned long long acc = 0;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
        longlong4 x;
        asm volatile("ld.global.nc.L1::no_allocate.v4.s64 {%0,%1,%2,%3}, [%4];" : "=l"(x.x), "=l"(x.y), "=l"(x.z), "=l"(x.w) : "l"(v + i));
        acc += x.x ^ x.y ^ x.z ^ x.w;
    }
    if (acc == 0x123456789ull) *out = acc;
}

__global__ void __launch_bounds__(256) copy_kernel(const longlong2 *v, longlong2 *o, int64_t n2)
{
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n2; i += (int64_t)gridDim.x * blockDim.x) o[i] = v[i];
}

struct LessThan {
    int64_t k;
    __host__ __device__ bool operator()(const int64_t &x) const { return x < k; }
};

struct PairLessThan {
    int64_t k;
    __host__ __device__ bool operator()(const thrust::tuple<uint32_t, int64_t> &x) const { return thrust::get<1>(x) < k; }
};

template <class F>
static float time_ms(F f, int reps)
{
    cudaEvent_t a, b;
    CK(cudaEventCreate(&a));
    CK(cudaEventCreate(&b));
    f();
    f();
    CK(cudaEventRecord(a));
    for (int i = 0; i < reps; i++) f();
    CK(cudaEventRecord(b));
    CK(cudaEventSynchronize(b));
    float ms = 0;
    CK(cudaEventElapsedTime(&ms, a, b));
    return ms / reps;
}

int main(int argc, char **argv)
{
    const int64_t n = argc > 1 ? atoll(argv[1]) : 1000000000LL;
    int64_t *in, *out;
    uint32_t *sel;
    unsigned long long *num;
    CK(cudaMalloc(&in, n * 8));
    CK(cudaMalloc(&out, n * 8));
    CK(cudaMalloc(&sel, n * 4));
    CK(cudaMalloc(&num, 8));
    gen<<<148 * 8, 256>>>(in, n);
    CK(cudaDeviceSynchronize());

    for (int per_sm : {4, 8, 16, 32}) {
        float ms = time_ms([&] { read_kernel<<<148 * per_sm, 256>>>((const longlong4 *)in, n / 4, num); }, 5);
        printf("{\"what\": \"read_ldg256\", \"ctas_per_sm\": %d, \"ms\": %.3f, \"gbs\": %.1f}\n", per_sm, ms, n * 8.0 / ms / 1e6);
    }
    {
        float ms = time_ms([&] { copy_kernel<<<148 * 16, 256>>>((const longlong2 *)in, (longlong2 *)out, n / 2); }, 5);
        printf("{\"what\": \"copy_kernel\", \"ms\": %.3f, \"gbs\": %.1f}\n", ms, n * 16.0 / ms / 1e6);
        ms = time_ms([&] { CK(cudaMemcpyAsync(out, in, n * 8, cudaMemcpyDeviceToDevice)); }, 5);
        printf("{\"what\": \"cudaMemcpy_d2d\", \"ms\": %.3f, \"gbs\": %.1f}\n", ms, n * 16.0 / ms / 1e6);
    }
    void *tmp = nullptr;
    size_t tmp_bytes = 0;
    for (double s : {0.01, 0.5, 0.99}) {
        LessThan pred{(int64_t)(s * 1099511627776.0)};
        size_t need = 0;
        CK(cub::DeviceSelect::If(nullptr, need, in, out, num, n, pred));
        if (need > tmp_bytes) { if (tmp) cudaFree(tmp); CK(cudaMalloc(&tmp, need)); tmp_bytes = need; }
        float ms = time_ms([&] { CK(cub::DeviceSelect::If(tmp, tmp_bytes, in, out, num, n, pred)); }, 5);
        unsigned long long kept = 0;
        CK(cudaMemcpy(&kept, num, 8, cudaMemcpyDeviceToHost));
        printf("{\"what\": \"cub_select_if_values\", \"sel\": %.2f, \"kept\": %llu, \"ms\": %.3f, \"gbs\": %.1f}\n", s, kept, ms,
               (n * 8.0 + kept * 8.0) / ms / 1e6);
    }
    for (double s : {0.01, 0.5, 0.99}) {
        PairLessThan pred{(int64_t)(s * 1099511627776.0)};
        auto zin = thrust::make_zip_iterator(thrust::make_tuple(thrust::counting_iterator<uint32_t>(0), in));
        auto zout = thrust::make_zip_iterator(thrust::make_tuple(sel, out));
        size_t need = 0;
        CK(cub::DeviceSelect::If(nullptr, need, zin, zout, num, n, pred));
        if (need > tmp_bytes) { if (tmp) cudaFree(tmp); CK(cudaMalloc(&tmp, need)); tmp_bytes = need; }
        float ms = time_ms([&] { CK(cub::DeviceSelect::If(tmp, tmp_bytes, zin, zout, num, n, pred)); }, 5);
        unsigned long long kept = 0;
        CK(cudaMemcpy(&kept, num, 8, cudaMemcpyDeviceToHost));
        printf("{\"what\": \"cub_select_if_sel_and_values\", \"sel\": %.2f, \"kept\": %llu, \"ms\": %.3f, \"gbs\": %.1f}\n", s, kept, ms,
               (n * 8.0 + kept * 12.0) / ms / 1e6);
    }
    return 0;
}
