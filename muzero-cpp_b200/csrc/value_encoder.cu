// value_encoder.cu — ValueEncoder::encode / decode (util.cpp:59-87) as device kernels behind the C ABI: the scalar <-> categorical
// support transforms of the value and reward heads.  decode is what the head epilogues fuse (heads.cuh, tower_fused.cuh); both
// directions are exported here for callers that hold logits or targets of their own (reanalyze targets, tests).
// Arithmetic follows the reference's fp32 tensor ops step by step; -fmad=false keeps every product / sum separately rounded.
#include <cmath>
#include <mutex>

#include "engine.h"

// Scratch of mzb_value_decode_static: the staged logits followed by the decoded scalars.  File scope, so every GPU that loads the
// library holds these 4 MiB until the process exits.  In host code the name is only a stand-in for the symbol: the device address
// comes from cudaGetSymbolAddress (static_scratch_address below), once per device with that device current.
#define MZB_STATIC_SCRATCH_FLOATS MZB_VALUE_DECODE_STATIC_MAX_FLOATS
static __device__ __align__(256) float g_static_scratch[MZB_STATIC_SCRATCH_FLOATS];

namespace mzb {
namespace {

// util.cpp:14-16 on a float tensor
__device__ __forceinline__ float contractive_f32(float x) {
    const float sg = (x > 0.f) ? 1.f : ((x < 0.f) ? -1.f : 0.f);
    return sg * (sqrtf(fabsf(x) + 1.f) - 1.f) + 0.001f * x;
}
// util.cpp:19-24
__device__ __forceinline__ float inverse_contractive_f32(float x) {
    const float eps = 0.001f;
    const float sg = (x > 0.f) ? 1.f : ((x < 0.f) ? -1.f : 0.f);
    const float inner = (sqrtf(1.f + 4.f * eps * (fabsf(x) + 1.f + eps)) - 1.f) / (2.f * eps);
    return sg * (inner * inner - 1.f);
}

// one thread per scalar: util.cpp:59-76.  support row = lower_mod at floor(steps), upper_mod at floor(steps) + 1, zero elsewhere
__global__ void value_encode_kernel(const float* __restrict__ x, int n, float* __restrict__ out, int S, float vmin, float vmax, float step, int contractive) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float red = contractive ? contractive_f32(x[i]) : x[i];
    red = fminf(fmaxf(red, vmin), vmax);
    const float num_steps = (red - vmin) / step;
    const float lower = floorf(num_steps);
    const float upper_mod = num_steps - lower;
    const float lower_mod = 1.0f - upper_mod;
    const long li = (long)lower;
    float* row = out + (long)i * S;
    for (int k = 0; k < S; ++k) row[k] = ((long)k == li ? 1.f : 0.f) * lower_mod + ((long)k == li + 1 ? 1.f : 0.f) * upper_mod;
}

// one warp per row: util.cpp:80-87 (softmax over the support, expected index, affine map, inverse contractive map)
__global__ void value_decode_kernel(const float* __restrict__ logits, int n, int S, int take_softmax, float* __restrict__ out, float vmin, float step, int contractive) {
    const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (row >= n) return;
    const float* lg = logits + (long)row * S;
    float mx = -3.402823466e38f;
    for (int k = lane; k < S; k += 32) mx = fmaxf(mx, lg[k]);
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    float sum = 0.f, ev = 0.f;
    for (int k = lane; k < S; k += 32) {
        const float p = take_softmax ? expf(lg[k] - mx) : lg[k];
        sum += p;
        ev += (float)k * p;
    }
    for (int o = 16; o > 0; o >>= 1) { sum += __shfl_xor_sync(0xffffffffu, sum, o); ev += __shfl_xor_sync(0xffffffffu, ev, o); }
    if (lane == 0) {
        const float e = take_softmax ? ev / sum : ev;
        const float v = e * step + vmin;
        out[row] = contractive ? inverse_contractive_f32(v) : v;
    }
}

int encoder_args(mzb_engine* e, int which, const void* a, const void* b, int n, const EncoderConsts** enc) {
    MZB_REQUIRE(e && a && b, "null argument");
    MZB_REQUIRE(which == 0 || which == 1, "which must be 0 (value encoder) or 1 (reward encoder)");
    MZB_REQUIRE(n >= 1, "n must be positive");
    *enc = which == 0 ? &e->venc : &e->renc;
    return MZB_OK;
}

// device address of g_static_scratch on `device` (which must be current): looked up once per device, then cached
int static_scratch_address(int device, float** out) {
    static std::mutex mu;
    static float* addr[64] = {};
    MZB_REQUIRE(device >= 0 && device < 64, "device ordinal %d outside [0, 64)", device);
    std::lock_guard<std::mutex> lk(mu);
    if (!addr[device]) {
        float* p = nullptr;
        MZB_CUDA(cudaGetSymbolAddress((void**)&p, g_static_scratch));
        MZB_REQUIRE(p != nullptr, "cudaGetSymbolAddress returned a null address for the static scratch");
        addr[device] = p;
    }
    *out = addr[device];
    return MZB_OK;
}

}  // namespace
}  // namespace mzb

using namespace mzb;

extern "C" {

int mzb_value_encode(mzb_engine* e, int which, const float* x, int n, float* support) {
    const EncoderConsts* enc = nullptr;
    int rc = encoder_args(e, which, x, support, n, &enc);
    if (rc) return rc;
    MZB_CUDA(cudaSetDevice(e->device));
    const int S = enc->support;
    DevBuf dx, ds;
    MZB_CUDA(dx.alloc((size_t)n * sizeof(float)));
    MZB_CUDA(ds.alloc((size_t)n * S * sizeof(float)));
    MZB_CUDA(cudaMemcpyAsync(dx.p, x, (size_t)n * sizeof(float), cudaMemcpyHostToDevice, e->stream));
    value_encode_kernel<<<(n + 127) / 128, 128, 0, e->stream>>>(dx.as<float>(), n, ds.as<float>(), S, (float)enc->min_value, (float)enc->max_value,
                                                               (float)enc->step, enc->contractive);
    MZB_CUDA(cudaGetLastError());
    MZB_CUDA(cudaMemcpyAsync(support, ds.p, (size_t)n * S * sizeof(float), cudaMemcpyDeviceToHost, e->stream));
    MZB_CUDA(cudaStreamSynchronize(e->stream));
    return MZB_OK;
}

int mzb_value_decode(mzb_engine* e, int which, const float* logits, int n, int take_softmax, float* x) {
    const EncoderConsts* enc = nullptr;
    int rc = encoder_args(e, which, logits, x, n, &enc);
    if (rc) return rc;
    MZB_CUDA(cudaSetDevice(e->device));
    const int S = enc->support;
    DevBuf dl, dx;
    MZB_CUDA(dl.alloc((size_t)n * S * sizeof(float)));
    MZB_CUDA(dx.alloc((size_t)n * sizeof(float)));
    MZB_CUDA(cudaMemcpyAsync(dl.p, logits, (size_t)n * S * sizeof(float), cudaMemcpyHostToDevice, e->stream));
    value_decode_kernel<<<(n * 32 + 127) / 128, 128, 0, e->stream>>>(dl.as<float>(), n, S, take_softmax, dx.as<float>(), (float)enc->min_value, (float)enc->step,
                                                                   enc->contractive);
    MZB_CUDA(cudaGetLastError());
    MZB_CUDA(cudaMemcpyAsync(x, dx.p, (size_t)n * sizeof(float), cudaMemcpyDeviceToHost, e->stream));
    MZB_CUDA(cudaStreamSynchronize(e->stream));
    return MZB_OK;
}

// mzb_value_decode without any allocation: the same kernel on the same stream, its staging in g_static_scratch instead of two
// cudaMalloc'ed buffers.  n * (support + 1) floats must fit MZB_STATIC_SCRATCH_FLOATS.  Nothing in the scratch has to start at
// zero: the copy overwrites the n * support logits and the kernel writes every one of the n outputs before they are read back;
// what a previous call left behind is never read.  All calls on a device share the one array: calls that overlap on the GPU
// (two engines on the same device driven from two threads) race — serialise them.
int mzb_value_decode_static(mzb_engine* e, int which, const float* logits, int n, int take_softmax, float* x) {
    const EncoderConsts* enc = nullptr;
    int rc = encoder_args(e, which, logits, x, n, &enc);
    if (rc) return rc;
    const int S = enc->support;
    MZB_REQUIRE((long long)n * (S + 1) <= (long long)MZB_STATIC_SCRATCH_FLOATS,
                "mzb_value_decode_static: %d rows of %d logits need %lld floats of scratch, the static array holds %d (use mzb_value_decode)", n, S,
                (long long)n * (S + 1), (int)MZB_STATIC_SCRATCH_FLOATS);
    MZB_CUDA(cudaSetDevice(e->device));
    float* scratch = nullptr;
    rc = static_scratch_address(e->device, &scratch);
    if (rc) return rc;
    float* dl = scratch;
    float* dx = scratch + (size_t)n * S;
    MZB_CUDA(cudaMemcpyAsync(dl, logits, (size_t)n * S * sizeof(float), cudaMemcpyHostToDevice, e->stream));
    value_decode_kernel<<<(n * 32 + 127) / 128, 128, 0, e->stream>>>(dl, n, S, take_softmax, dx, (float)enc->min_value, (float)enc->step, enc->contractive);
    MZB_CUDA(cudaGetLastError());
    MZB_CUDA(cudaMemcpyAsync(x, dx, (size_t)n * sizeof(float), cudaMemcpyDeviceToHost, e->stream));
    MZB_CUDA(cudaStreamSynchronize(e->stream));
    return MZB_OK;
}

}  // extern "C"
