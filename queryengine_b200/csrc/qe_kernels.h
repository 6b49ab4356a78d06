// qe_kernels.h -- host-callable launchers of the ahead-of-time compiled kernels
// in qe_kernels.cu.
#ifndef QE_KERNELS_H
#define QE_KERNELS_H

#include <cstdint>
#include <cuda_runtime.h>

#define QE_INTERP_STEPS 1
#define QE_INTERP_VEC 2

struct QeParams;

namespace qe {

// accumulator words of a group-by (GroupPlan of qe_codegen.h), by value into the helper kernels
struct GroupWords {
    int nwords;
    unsigned char kind[48];
    unsigned long long init[48];
};

// mode: 0 filter, 1 project, 2 aggregate (== KernelMode)
cudaError_t launch_interp(int mode, const QeParams &p, int grid, cudaStream_t stream);
int interp_blocks_per_sm(int mode);
// folds [nsub][nagg][4] sub-tile partials into [ceil(nsub / 64)][nagg][4] super-tile partials (fixed tree per super-tile)
cudaError_t launch_agg_final(const double *sub_partials, int64_t nsub, int nagg, double *super_partials, cudaStream_t stream);
// group-by helpers: ctl = [0] groups claimed [1] overflow flags [2] records collected [3] spare
cudaError_t launch_groupby_init(uint32_t *state, uint64_t *acc, uint32_t *ctl, int64_t nslots, const GroupWords &gw, cudaStream_t stream);
cudaError_t launch_groupby_merge(const uint64_t *part, const uint32_t *part_slot, uint64_t *acc, int nctas, int lmax,
                                 const GroupWords &gw, cudaStream_t stream);
cudaError_t launch_groupby_collect(const uint32_t *state, const uint64_t *keys, const uint32_t *nulls, const uint64_t *acc,
                                   int64_t nslots, int nkeys, int nwords, uint64_t *records, uint32_t *ctl, uint32_t capacity,
                                   cudaStream_t stream);
// *dst = *src by one thread; dst may be pinned host memory (mapped under unified addressing).  For the counts a stream reads
// back per chunk: a tiny cudaMemcpyAsync would queue behind the large downloads on the copy engine and hold the compute
// stream for milliseconds.
cudaError_t launch_store_u64(unsigned long long *dst, const unsigned long long *src, cudaStream_t stream);
// kind: 0 read `bytes` of src, 1 copy src -> dst, 2 write dst
cudaError_t launch_yardstick(const void *src, void *dst, int64_t bytes, int kind, int sms, void *sink, cudaStream_t stream);
cudaError_t launch_generate(void *values, uint32_t *validity, int type, int kind, uint64_t seed, double p0, double p1,
                            double null_rate, int64_t row_offset, int64_t nrows, cudaStream_t stream);

}  // namespace qe
#endif
