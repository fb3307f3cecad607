// lensrt — the fast-arithmetic flavour of the spot kernels (lrt_spot_kernel.cuh compiled with LRT_FAST_ARITH): lens-design
// loops do not need the reference's last-ulp rounding, only its algorithm. Selected by lrt_spot_params.fast.
#define LRT_FAST_ARITH 1
#define LRT_SPOT_FLAVOUR 1
#define LRT_SPOT_LAUNCH launch_spot_flavour_fast
#include "lrt_spot_kernel.cuh"

namespace lrt {
cudaError_t launch_spot_fast(const SpotArgs &A, bool adjoint, int grid, cudaStream_t stream) { return launch_spot_flavour_fast(A, adjoint, grid, stream); }
}
