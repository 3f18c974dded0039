// net_fp32.cu — MZB_FP32_CHECK backend: the direct executor with fp32 activation storage (see net_direct.cuh).
#include "net_direct.cuh"

namespace mzb {

__global__ void initial_reward_kernel(double* reward, int B, int support, float step, float min_value, int contractive) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= B) return;
    float ex = (float)(support / 2);
    float v = ex * step + min_value;
    if (contractive) {
        const float eps = 0.001f;
        float s = (v > 0.f) ? 1.f : ((v < 0.f) ? -1.f : 0.f);
        float inner = sqrtf(1.f + (4.f * eps) * (fabsf(v) + 1.f + eps)) - 1.f;
        float q = inner / (2.f * eps);
        v = s * (q * q - 1.f);
    }
    reward[i] = (double)v;
}

NetBackend* make_fp32_backend() { return new direct::DirectBackend<float>(); }

}  // namespace mzb
