// heads.cuh — the three small heads behind the towers, one kernel each launch:
//   conv1x1(C -> red, bias)  ->  flatten channel-major (r*HW + p)  ->  MLP  ->  output transform
// fused with the policy softmax (vprnet.cpp:118-119,178-179) or the support-to-scalar decode
// (ValueEncoder::decode, util.cpp:80-87 + inverse contractive map util.cpp:19-24).
// < 0.3 % of the network FLOPs: CUDA-core fp32 math for both precision modes.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>

namespace mzb {

struct HeadDev {
    int C, red, hw, n_layers;
    int sizes[10];          // sizes[0] = red*hw, ..., sizes[n_layers] = out
    const float* w1;        // [red][C]
    const float* b1;        // [red]
    const float* W[9];      // [out][in] per layer
    const float* b[9];
    int out_kind;           // 0 = policy softmax, 1 = support -> scalar
    // decode constants (float, as the reference computes in float32)
    float step, min_value;
    int contractive;
};

__device__ __forceinline__ float head_ld(const float* p) { return *p; }
__device__ __forceinline__ float head_ld(const __nv_bfloat16* p) { return __bfloat162float(*p); }

// x: NHWC activations, sample b at x + slot(b) * sample_stride.  One CTA (128 threads) per sample.
// dynamic smem: float xs[hw*C] + float bufA[maxw] + float bufB[maxw]   (maxw = max layer width incl. red*hw)
template <class T>
__global__ void __launch_bounds__(128) head_kernel(const T* __restrict__ x, const int* __restrict__ slots,
                                                   long sample_stride, HeadDev hd, double* __restrict__ out,
                                                   int out_stride, int maxw) {
    extern __shared__ float smem[];
    const int b = blockIdx.x;
    const int tid = threadIdx.x;
    const int lane = tid & 31, warp = tid >> 5;
    float* xs = smem;
    float* bufA = xs + hd.hw * hd.C;
    float* bufB = bufA + maxw;
    const T* xb = x + (long)(slots ? slots[b] : b) * sample_stride;
    for (int i = tid; i < hd.hw * hd.C; i += 128) xs[i] = head_ld(xb + i);
    __syncthreads();
    // conv1x1: out[r][p] = b1[r] + sum_c w1[r][c] * x[p][c]; one warp per output, lanes stride over c
    const int n0 = hd.red * hd.hw;
    for (int o = warp; o < n0; o += 4) {
        int r = o / hd.hw, p = o - r * hd.hw;
        double acc = 0.0;   // fp64 accumulation: the heads are tiny and feed the error-amplifying decode
        for (int c = lane; c < hd.C; c += 32) acc += (double)(hd.w1[r * hd.C + c] * xs[p * hd.C + c]);
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, off);
        if (lane == 0) bufA[o] = (float)(acc + (double)hd.b1[r]);
    }
    __syncthreads();
    float* cur = bufA;
    float* nxt = bufB;
    for (int l = 0; l < hd.n_layers; ++l) {
        const int nin = hd.sizes[l], nout = hd.sizes[l + 1];
        const float* W = hd.W[l];
        const float* bias = hd.b[l];
        for (int j = warp; j < nout; j += 4) {
            double dacc = 0.0;
            for (int i = lane; i < nin; i += 32) dacc += (double)(W[(long)j * nin + i] * cur[i]);
#pragma unroll
            for (int off = 16; off > 0; off >>= 1) dacc += __shfl_xor_sync(0xffffffffu, dacc, off);
            if (lane == 0) {
                float acc = (float)(dacc + (double)bias[j]);
                if (l < hd.n_layers - 1) acc = fmaxf(acc, 0.f);
                nxt[j] = acc;
            }
        }
        __syncthreads();
        float* t = cur; cur = nxt; nxt = t;
    }
    if (tid == 0) {
        const int n = hd.sizes[hd.n_layers];
        float mx = cur[0];
        for (int i = 1; i < n; ++i) mx = fmaxf(mx, cur[i]);
        float sum = 0.f;
        for (int i = 0; i < n; ++i) { float e = expf(cur[i] - mx); nxt[i] = e; sum += e; }
        if (hd.out_kind == 0) {
            for (int i = 0; i < n; ++i) out[(long)b * out_stride + i] = (double)(nxt[i] / sum);
        } else {
            float ex = 0.f;
            for (int i = 0; i < n; ++i) ex += (float)i * (nxt[i] / sum);
            float v = ex * hd.step + hd.min_value;
            if (hd.contractive) {
                const float eps = 0.001f;
                float s = (v > 0.f) ? 1.f : ((v < 0.f) ? -1.f : 0.f);
                float inner = sqrtf(1.f + (4.f * eps) * (fabsf(v) + 1.f + eps)) - 1.f;
                float q = inner / (2.f * eps);
                v = s * (q * q - 1.f);
            }
            out[(long)b * out_stride] = (double)v;
        }
    }
}

// InitialInference's reward: log(one_hot(S/2)) decodes to the support centre (models.cpp:252-256).
// Computed with the same float formula so that a non-symmetric support still matches.
__global__ void initial_reward_kernel(double* reward, int B, int support, float step, float min_value, int contractive);

}  // namespace mzb
