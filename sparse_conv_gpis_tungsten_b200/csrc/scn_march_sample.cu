// scn_march_sample.cu — the k_march instantiations behind scn_sample_distance_batch (3D media); see scn_march.cuh
#include <cuda_runtime.h>

#include "scn_device.cuh"
#include "scn_fast.cuh"
#include "scn_march.cuh"

namespace scn {
cudaError_t march_launch_sample(const Params &P, const MarchArgs &A, int block, int sm_count, cudaStream_t s) { return launch_march_t<false>(P, A, block, sm_count, s); }
SCN_DEFINE_BOUNDS_COLLECT(bounds_collect_march_sample)
} // namespace scn
