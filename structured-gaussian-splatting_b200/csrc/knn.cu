// knn.cu - distCUDA2 (mean squared distance to the 3 nearest neighbours).  Placeholder until the
// kernels land: reports GSR_EINVAL so nothing can silently succeed.
#include "gsr_common.cuh"
namespace gsr {
int launch_dist2(int, const float*, float*, GsrAlloc, cudaStream_t) { return GSR_EINVAL; }
}  // namespace gsr
