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

constexpr int HEAD_GS = 8;        // samples per CTA: every MLP weight fetched from L2 is used 8 times
constexpr int HEAD_THREADS = 256;

template <class T> struct HeadAcc { typedef float type; };
template <> struct HeadAcc<float> { typedef double type; };   // check mode: fp64 accumulation (feeds the error-amplifying decode)

inline size_t head_smem_bytes(const HeadDev& hd, int maxw, int maxw2) {
    return ((size_t)hd.hw * (hd.C + 1) + (size_t)hd.red * hd.C + (size_t)HEAD_GS * (maxw + maxw2)) * sizeof(float);
}

// x: NHWC activations, sample b at x + slot(b) * sample_stride.  One CTA per HEAD_GS samples.
// dynamic smem: xs[hw][C+1] | w1s[red][C] | bufA[GS][maxw] | bufB[GS][maxw2]     (maxw = red*hw, maxw2 = widest later layer)
template <class T>
__global__ void __launch_bounds__(HEAD_THREADS) head_kernel(const T* __restrict__ x, const int* __restrict__ slots,
                                                            long sample_stride, int B, HeadDev hd, double* __restrict__ out,
                                                            int out_stride, int maxw, int maxw2) {
    typedef typename HeadAcc<T>::type acc_t;
    extern __shared__ float smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int b0 = blockIdx.x * HEAD_GS;
    const int C = hd.C, hw = hd.hw, n0 = hd.red * hd.hw;
    float* xs = smem;
    float* w1s = xs + hw * (C + 1);
    float* bufA = w1s + hd.red * C;
    float* bufB = bufA + HEAD_GS * maxw;
    for (int i = tid; i < hd.red * C; i += HEAD_THREADS) w1s[i] = hd.w1[i];
    // conv1x1 + channel-major flatten, one sample at a time: flat[g][r*hw + p] = b1[r] + sum_c w1[r][c] * x[p][c]
    for (int g = 0; g < HEAD_GS; ++g) {
        const int b = b0 + g;
        __syncthreads();
        if (b < B) {
            const T* xb = x + (long)(slots ? slots[b] : b) * sample_stride;
            for (int i = tid; i < hw * C; i += HEAD_THREADS) { const int p = i / C, c = i - p * C; xs[p * (C + 1) + c] = head_ld(xb + i); }
        }
        __syncthreads();
        for (int o = tid; o < n0; o += HEAD_THREADS) {
            const int r = o / hw, p = o - r * hw;
            acc_t acc = 0;
            if (b < B) {
                const float* wr = w1s + r * C;
                const float* xr = xs + p * (C + 1);
#pragma unroll 8
                for (int c = 0; c < C; ++c) acc += (acc_t)(wr[c] * xr[c]);
            }
            bufA[g * maxw + o] = (float)(acc + (acc_t)hd.b1[r]);
        }
    }
    __syncthreads();
    float* cur = bufA; int cur_w = maxw;
    float* nxt = bufB; int nxt_w = maxw2;
    for (int l = 0; l < hd.n_layers; ++l) {
        const int nin = hd.sizes[l], nout = hd.sizes[l + 1];
        const float* __restrict__ W = hd.W[l];
        const float* __restrict__ bias = hd.b[l];
        for (int j = warp; j < nout; j += HEAD_THREADS / 32) {
            acc_t acc[HEAD_GS];
#pragma unroll
            for (int g = 0; g < HEAD_GS; ++g) acc[g] = 0;
            for (int i = lane; i < nin; i += 32) {
                const float w = W[(long)j * nin + i];
#pragma unroll
                for (int g = 0; g < HEAD_GS; ++g) acc[g] += (acc_t)(w * cur[g * cur_w + i]);
            }
#pragma unroll
            for (int g = 0; g < HEAD_GS; ++g) {
#pragma unroll
                for (int off = 16; off > 0; off >>= 1) acc[g] += __shfl_xor_sync(0xffffffffu, acc[g], off);
            }
            if (lane == 0) {
#pragma unroll
                for (int g = 0; g < HEAD_GS; ++g) {
                    float v = (float)(acc[g] + (acc_t)bias[j]);
                    if (l < hd.n_layers - 1) v = fmaxf(v, 0.f);
                    nxt[g * nxt_w + j] = v;
                }
            }
        }
        __syncthreads();
        float* t = cur; cur = nxt; nxt = t;
        const int tw = cur_w; cur_w = nxt_w; nxt_w = tw;
    }
    if (tid < HEAD_GS && b0 + tid < B) {
        const int b = b0 + tid;
        const float* lg = cur + tid * cur_w;
        float* ex = nxt + tid * nxt_w;
        const int n = hd.sizes[hd.n_layers];
        float mx = lg[0];
        for (int i = 1; i < n; ++i) mx = fmaxf(mx, lg[i]);
        float sum = 0.f;
        for (int i = 0; i < n; ++i) { float e = expf(lg[i] - mx); ex[i] = e; sum += e; }
        if (hd.out_kind == 0) {
            for (int i = 0; i < n; ++i) out[(long)b * out_stride + i] = (double)(ex[i] / sum);
        } else {
            float ev = 0.f;
            for (int i = 0; i < n; ++i) ev += (float)i * (ex[i] / sum);
            float v = ev * hd.step + hd.min_value;
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
