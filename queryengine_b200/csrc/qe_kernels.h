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

// mode: 0 filter, 1 project, 2 aggregate (== KernelMode)
cudaError_t launch_interp(int mode, const QeParams &p, int grid, cudaStream_t stream);
int interp_blocks_per_sm(int mode);
cudaError_t launch_agg_final(const double *partials, int nblocks, int nagg, double *out, cudaStream_t stream);
cudaError_t launch_generate(void *values, uint32_t *validity, int type, int kind, uint64_t seed, double p0, double p1,
                            double null_rate, int64_t row_offset, int64_t nrows, cudaStream_t stream);

}  // namespace qe
#endif
