// placeholder — replaced by the tcgen05 kernel
#include "common.cuh"
#include "kernels.h"
namespace cslam {
bool k2_supported(int, int) { return false; }
int k2_grid(long long, int num_sms) { return num_sms; }
cudaError_t k2_prepare(const float*, int, int, int, float*, float*, cudaStream_t) { return cudaErrorNotSupported; }
cudaError_t k2_launch(const K2Args&, int, cudaStream_t, const char** err) { if (err) *err = "k2 not built"; return cudaErrorNotSupported; }
}
