// scn_march1d.cu — the k_cond1d / k_march1d / k_finish1d instantiations (1D-sampling media); see scn_march1d.cuh
#include <cuda_runtime.h>

#include "scn_device.cuh"
#include "scn_fast.cuh"
#define SCN_MARCH1D_IMPL 1
#include "scn_march1d.cuh"

namespace scn {
cudaError_t march1d_launch_sample(const Params &P, const MarchArgs &A, int sm_count, cudaMemPool_t pool, cudaStream_t s) { return launch_march1d_t<false>(P, A, sm_count, pool, s); }
cudaError_t march1d_launch_transmit(const Params &P, const MarchArgs &A, int sm_count, cudaMemPool_t pool, cudaStream_t s) { return launch_march1d_t<true>(P, A, sm_count, pool, s); }
cudaError_t march1d_launch_sample_static(const Params &P, const MarchArgs &A, int sm_count, void *scratch, cudaStream_t s) { return launch_march1d_sample_on(P, A, sm_count, scratch, s); }
SCN_DEFINE_BOUNDS_COLLECT(bounds_collect_march1d)
} // namespace scn
