// net_direct.cuh — direct (CUDA-core) executor of the VPRNet towers, templated on the activation storage type.
//   T = float          : MZB_FP32_CHECK, the check mode whose only job is to agree with libtorch fp32 to <= 1e-4
//                        (north star): an on-device reference for the tensor-core path and a high-precision C-ABI mode.
//   T = __nv_bfloat16  : base class of the bf16 backend (net_bf16.cu), which overrides fast_conv() to run every
//                        eligible 3x3 convolution on tcgen05; the layers that stay here are the < 1 % odd shapes.
// Activations are NHWC; BN is pre-folded; accumulation is fp32 per 16-channel chunk folded into fp64.
// Replaces (in this mode): model_layers.cpp:70-162 forward passes, models.cpp:16-26 normalisation,
// models.cpp:105-117,153-166 heads, vprnet.cpp:103-127 output conversion.
#pragma once
#include <cfloat>
#include <cstring>

#include "heads.cuh"
#include "net_backend.h"

namespace mzb {

namespace direct {

constexpr int TILE = 8;    // output tile 8x8 pixels
constexpr int TC = 64;     // couts per CTA
constexpr int CK = 16;     // input channels per smem chunk

__device__ __forceinline__ float ldf(const float* p) { return *p; }
__device__ __forceinline__ float ldf(const __nv_bfloat16* p) { return __bfloat162float(*p); }
__device__ __forceinline__ void stf(float* p, float v) { *p = v; }
__device__ __forceinline__ void stf(__nv_bfloat16* p, float v) { *p = __float2bfloat16(v); }

template <class T>
struct ConvArgs {
    const T* in; const int* in_slot; long in_stride;      // NHWC [in_h][in_w][cin] per sample
    T* out; const int* out_slot; long out_stride;         // NHWC [out_h][out_w][cout]
    const T* res; const int* res_slot; long res_stride;   // optional residual, same shape as out
    const float* w;                                           // [9][cin][cout]
    const float* bias;                                        // [cout]
    const float* act_table; const int* action;                // [A][out_hw][cout], [B]
    int cin, cout, stride, in_h, in_w, out_h, out_w, relu;
};

// y[b][oy][ox][co] = act( sum_{ky,kx,ci} x[b][oy*s+ky-1][ox*s+kx-1][ci] * w[ky*3+kx][ci][co] + bias[co] (+table) (+res) )
template <class T>
__global__ void __launch_bounds__(256) conv3_direct_kernel(ConvArgs<T> a) {
    extern __shared__ float smem[];
    const int s = a.stride;
    const int PW = TILE * s + 2;                 // patch width/height
    float* patch = smem;                          // [CK][PW][PW]
    float* wsm = smem + CK * PW * PW;             // [9][CK][TC]
    const int tiles_x = (a.out_w + TILE - 1) / TILE;
    const int ty = blockIdx.x / tiles_x, tx = blockIdx.x % tiles_x;
    const int co0 = blockIdx.y * TC;
    const int b = blockIdx.z;
    const int tid = threadIdx.x;
    const int pix = tid & 63, cg = tid >> 6;
    const int py = pix >> 3, px = pix & 7;
    const T* inb = a.in + (long)(a.in_slot ? a.in_slot[b] : b) * a.in_stride;
    // fp32 partial sums per 16-channel chunk (144 terms), folded into fp64 totals: keeps the check mode's rounding
    // error an order of magnitude below the fp32 reference's own, so the 1e-4 bar is met with margin
    double dacc[16];
    float acc[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) dacc[j] = 0.0;
    const int iy0 = ty * TILE * s - 1, ix0 = tx * TILE * s - 1;
    for (int c0 = 0; c0 < a.cin; c0 += CK) {
#pragma unroll
        for (int j = 0; j < 16; ++j) acc[j] = 0.f;
        // stage the input patch (channel-major so lanes read consecutive x)
        for (int i = tid; i < CK * PW * PW; i += 256) {
            int ci = i % CK;
            int r = i / CK;
            int xx = r % PW, yy = r / PW;
            int iy = iy0 + yy, ix = ix0 + xx, c = c0 + ci;
            float v = 0.f;
            if (iy >= 0 && iy < a.in_h && ix >= 0 && ix < a.in_w && c < a.cin) v = ldf(inb + ((long)iy * a.in_w + ix) * a.cin + c);
            patch[(ci * PW + yy) * PW + xx] = v;
        }
        for (int i = tid; i < 9 * CK * TC; i += 256) {
            int co = i % TC;
            int r = i / TC;
            int ci = r % CK, tap = r / CK;
            int c = c0 + ci;
            float v = 0.f;
            if (c < a.cin && co0 + co < a.cout) v = a.w[((long)tap * a.cin + c) * a.cout + co0 + co];
            wsm[i] = v;
        }
        __syncthreads();
#pragma unroll
        for (int tap = 0; tap < 9; ++tap) {
            const int ky = tap / 3, kx = tap % 3;
            const float* prow = patch + (py * s + ky) * PW + (px * s + kx);
            const float* wrow = wsm + tap * CK * TC + cg * 16;
#pragma unroll 4
            for (int ci = 0; ci < CK; ++ci) {
                float xv = prow[ci * PW * PW];
                const float4* w4 = reinterpret_cast<const float4*>(wrow + ci * TC);
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    float4 wv = w4[q];
                    acc[q * 4 + 0] += xv * wv.x; acc[q * 4 + 1] += xv * wv.y;
                    acc[q * 4 + 2] += xv * wv.z; acc[q * 4 + 3] += xv * wv.w;
                }
            }
        }
        __syncthreads();
#pragma unroll
        for (int j = 0; j < 16; ++j) dacc[j] += (double)acc[j];
    }
    const int oy = ty * TILE + py, ox = tx * TILE + px;
    if (oy >= a.out_h || ox >= a.out_w) return;
    const int opix = oy * a.out_w + ox;
    T* outb = a.out + (long)(a.out_slot ? a.out_slot[b] : b) * a.out_stride + (long)opix * a.cout;
    const T* resb = a.res ? a.res + (long)(a.res_slot ? a.res_slot[b] : b) * a.res_stride + (long)opix * a.cout : nullptr;
    const float* tab = a.act_table ? a.act_table + ((long)a.action[b] * a.out_h * a.out_w + opix) * a.cout : nullptr;
#pragma unroll
    for (int j = 0; j < 16; ++j) {
        int co = co0 + cg * 16 + j;
        if (co >= a.cout) break;
        double dv = dacc[j] + (double)a.bias[co];
        if (tab) dv += (double)tab[co];
        if (resb) dv += (double)ldf(resb + co);
        float v = (float)dv;
        if (a.relu) v = fmaxf(v, 0.f);
        stf(outb + co, v);
    }
}

// AvgPool2d(3, stride 2, padding 1), count_include_pad=true -> divisor always 9 (model_layers.cpp:31-33)
template <class T>
__global__ void avgpool3_kernel(const T* in, T* out, int B, int C, int in_h, int in_w, int out_h, int out_w) {
    long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    long total = (long)B * out_h * out_w * C;
    if (i >= total) return;
    int c = i % C;
    long r = i / C;
    int ox = r % out_w; r /= out_w;
    int oy = r % out_h;
    int b = r / out_h;
    float acc = 0.f;
    for (int ky = 0; ky < 3; ++ky)
        for (int kx = 0; kx < 3; ++kx) {
            int iy = oy * 2 + ky - 1, ix = ox * 2 + kx - 1;
            if (iy >= 0 && iy < in_h && ix >= 0 && ix < in_w) acc += ldf(in + (((long)b * in_h + iy) * in_w + ix) * C + c);
        }
    stf(out + i, acc / 9.0f);
}

// normalize_hidden_state (models.cpp:16-26): per (sample, channel) min/max over pixels; scale < 1e-5 -> 1e-5.
// normalize == 0 turns it into a plain copy into the pool slot.
template <class T>
__global__ void normalize_kernel(const T* in, T* pool, const int* out_slot, long slot_stride, int hw, int C, int normalize) {
    const int b = blockIdx.x;
    const T* xb = in + (long)b * hw * C;
    T* ob = pool + (long)(out_slot ? out_slot[b] : b) * slot_stride;
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
        float mn = FLT_MAX, mx = -FLT_MAX;
        for (int p = 0; p < hw; ++p) { float v = ldf(xb + p * C + c); mn = fminf(mn, v); mx = fmaxf(mx, v); }
        float scale = mx - mn;
        if (scale < 1e-5f) scale = 1e-5f;
        for (int p = 0; p < hw; ++p) {
            float v = ldf(xb + p * C + c);
            stf(ob + p * C + c, normalize ? (v - mn) / scale : v);
        }
    }
}

// rows [B][C][HW] (NCHW flat, the reference's Observation order) <-> NHWC slots
template <class T>
__global__ void nchw_to_nhwc_kernel(const float* in, T* out, const int* out_slot, long out_stride, int B, int C, int hw) {
    long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    long total = (long)B * C * hw;
    if (i >= total) return;
    int c = i % C;
    long r = i / C;
    int p = r % hw;
    int b = r / hw;
    stf(out + (long)(out_slot ? out_slot[b] : b) * out_stride + (long)p * C + c, in[((long)b * C + c) * hw + p]);
}
template <class T>
__global__ void nhwc_to_nchw_kernel(const T* in, const int* in_slot, long in_stride, float* out, int B, int C, int hw) {
    long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    long total = (long)B * C * hw;
    if (i >= total) return;
    int p = i % hw;
    long r = i / hw;
    int c = r % C;
    int b = r / C;
    out[i] = ldf(in + (long)(in_slot ? in_slot[b] : b) * in_stride + (long)p * C + c);
}

struct ConvDev {
    size_t w_off = 0, bias_off = 0, table_off = 0;
    bool has_table = false;
    int cin, cout, stride, in_h, in_w, out_h, out_w, relu;
    int tower = 0, index = 0;   // position in the plan (for backends that keep extra per-conv state)
};
struct HeadOff {
    size_t w1, b1;
    size_t W[9], b[9];
};

template <class T>
__global__ void copy_slots_kernel(const T* in, const int* slot, long stride, T* out, long elems, long total) {
    long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= total) return;
    long b = i / elems, r = i - b * elems;
    out[i] = in[(long)slot[b] * stride + r];
}

// One conv launch request, as the tower executor sees it (all buffers hold T, NHWC).
template <class T>
struct ConvCall {
    const ConvDev* conv;
    int B;
    const T* in; const int* in_slot; long in_stride;
    T* out; const int* out_slot; long out_stride;
    const T* res; const int* res_slot; long res_stride;
    const int* action;
};

template <class T>
class DirectBackend : public NetBackend {
public:
    const char* name() const override { return sizeof(T) == 4 ? "fp32-check" : "bf16-direct"; }

    int upload(const NetPlan& plan) override {
        plan_ = &plan;
        plan_dims_ = plan.dims;
        normalize_ = plan.normalize;
        towers_[0] = &plan.representation; towers_[1] = &plan.dynamics; towers_[2] = &plan.prediction;
        std::vector<float> blob;
        for (int t = 0; t < 3; ++t) {
            steps_[t] = towers_[t]->steps;
            convs_[t].clear();
            int index = 0;
            for (const Conv3Host& c : towers_[t]->convs) {
                ConvDev d;
                d.cin = c.cin; d.cout = c.cout; d.stride = c.stride; d.in_h = c.in_h; d.in_w = c.in_w;
                d.out_h = c.out_h; d.out_w = c.out_w; d.relu = c.relu;
                d.tower = t; d.index = index++;
                d.w_off = blob.size();
                blob.resize(blob.size() + (size_t)9 * c.cin * c.cout);
                for (int o = 0; o < c.cout; ++o)
                    for (int tp = 0; tp < 9; ++tp)
                        for (int ci = 0; ci < c.cin; ++ci)
                            blob[d.w_off + ((size_t)tp * c.cin + ci) * c.cout + o] = c.w[((size_t)o * 9 + tp) * c.cin + ci];
                d.bias_off = blob.size();
                blob.insert(blob.end(), c.bias.begin(), c.bias.end());
                d.has_table = !c.act_table.empty();
                d.table_off = blob.size();
                blob.insert(blob.end(), c.act_table.begin(), c.act_table.end());
                while (blob.size() % 4) blob.push_back(0.f);
                convs_[t].push_back(d);
            }
            max_act_[t] = towers_[t]->max_act_elems;
        }
        const HeadHost* hh[3] = {&plan.reward_head, &plan.policy_head, &plan.value_head};
        for (int h = 0; h < 3; ++h) {
            heads_[h] = HeadDev();
            heads_[h].C = hh[h]->C; heads_[h].red = hh[h]->red; heads_[h].hw = hh[h]->hw;
            heads_[h].n_layers = (int)hh[h]->W.size();
            for (size_t i = 0; i < hh[h]->sizes.size(); ++i) heads_[h].sizes[i] = hh[h]->sizes[i];
            hoff_[h].w1 = blob.size(); blob.insert(blob.end(), hh[h]->w1.begin(), hh[h]->w1.end());
            hoff_[h].b1 = blob.size(); blob.insert(blob.end(), hh[h]->b1.begin(), hh[h]->b1.end());
            for (int l = 0; l < heads_[h].n_layers; ++l) {
                hoff_[h].W[l] = blob.size(); blob.insert(blob.end(), hh[h]->W[l].begin(), hh[h]->W[l].end());
                hoff_[h].b[l] = blob.size(); blob.insert(blob.end(), hh[h]->b[l].begin(), hh[h]->b[l].end());
            }
            int mw2 = 1;
            for (size_t i = 1; i < hh[h]->sizes.size(); ++i) mw2 = hh[h]->sizes[i] > mw2 ? hh[h]->sizes[i] : mw2;
            head_maxw_[h] = hh[h]->sizes[0];
            head_maxw2_[h] = mw2;
        }
        heads_[0].out_kind = 1; heads_[1].out_kind = 0; heads_[2].out_kind = 1;
        heads_[0].step = (float)plan.reward_enc.step; heads_[0].min_value = (float)plan.reward_enc.min_value; heads_[0].contractive = plan.reward_enc.contractive;
        heads_[2].step = (float)plan.value_enc.step; heads_[2].min_value = (float)plan.value_enc.min_value; heads_[2].contractive = plan.value_enc.contractive;
        renc_ = plan.reward_enc;
        // backend-specific extra weights (e.g. bf16 tensor-core packing) are appended to the same blob so that one
        // broadcast moves everything
        while (blob.size() % 32) blob.push_back(0.f);   // extras start 128-byte aligned (bulk copies need 16)
        std::vector<uint8_t> extra;
        int rc = pack_extra(plan, &extra);
        if (rc) return rc;
        fp32_bytes_ = blob.size() * sizeof(float);
        size_t total = fp32_bytes_ + extra.size();
        if (blob_.bytes != total) MZB_CUDA(blob_.alloc(total));
        MZB_CUDA(cudaMemcpy(blob_.p, blob.data(), fp32_bytes_, cudaMemcpyHostToDevice));
        if (!extra.empty()) MZB_CUDA(cudaMemcpy((uint8_t*)blob_.p + fp32_bytes_, extra.data(), extra.size(), cudaMemcpyHostToDevice));
        const float* base = blob_.as<float>();
        for (int h = 0; h < 3; ++h) {
            heads_[h].w1 = base + hoff_[h].w1; heads_[h].b1 = base + hoff_[h].b1;
            for (int l = 0; l < heads_[h].n_layers; ++l) { heads_[h].W[l] = base + hoff_[h].W[l]; heads_[h].b[l] = base + hoff_[h].b[l]; }
        }
        // opt in to large dynamic smem once
        int max_smem = 0;
        for (int t = 0; t < 3; ++t) for (auto& c : convs_[t]) { int sm = conv_smem(c.stride); if (sm > max_smem) max_smem = sm; }
        MZB_CUDA(cudaFuncSetAttribute(conv3_direct_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem));
        int hs = 0;
        for (int h = 0; h < 3; ++h) { int sm = head_smem(h); if (sm > hs) hs = sm; }
        MZB_CUDA(cudaFuncSetAttribute(head_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, hs));
        uploaded_ = true;
        return MZB_OK;
    }
    void* blob_ptr() override { return blob_.p; }
    size_t blob_bytes() const override { return blob_.bytes; }

    int alloc_pool(size_t n_slots, int max_batch) override {
        MZB_REQUIRE(uploaded_, "alloc_pool before upload");
        slot_elems_ = (size_t)plan_dims_.hidden_size();
        MZB_CUDA(pool_.alloc(n_slots * slot_elems_ * sizeof(T)));
        MZB_CUDA(cudaMemset(pool_.p, 0, pool_.bytes));
        max_batch_ = max_batch;
        size_t mx = 0;
        for (int t = 0; t < 3; ++t) if ((size_t)max_act_[t] > mx) mx = max_act_[t];
        mx = act_elems(mx);
        MZB_CUDA(bufA_.alloc((size_t)max_batch * mx * sizeof(T)));
        MZB_CUDA(bufB_.alloc((size_t)max_batch * mx * sizeof(T)));
        MZB_CUDA(bufC_.alloc((size_t)max_batch * mx * sizeof(T)));
        return MZB_OK;
    }
    size_t pool_bytes() const override { return pool_.bytes; }
    size_t slot_bytes() const override { return slot_elems_ * sizeof(T); }

    int import_hidden(const float* d_nchw, int B, const int* d_slot, cudaStream_t s) override {
        long total = (long)B * slot_elems_;
        nchw_to_nhwc_kernel<T><<<(unsigned)((total + 255) / 256), 256, 0, s>>>(d_nchw, pool_.as<T>(), d_slot, (long)slot_elems_, B, plan_dims_.C, plan_dims_.hw());
        MZB_CUDA(cudaGetLastError());
        return MZB_OK;
    }
    int export_hidden(float* d_nchw, int B, const int* d_slot, cudaStream_t s) override {
        long total = (long)B * slot_elems_;
        nhwc_to_nchw_kernel<T><<<(unsigned)((total + 255) / 256), 256, 0, s>>>(pool_.as<T>(), d_slot, (long)slot_elems_, d_nchw, B, plan_dims_.C, plan_dims_.hw());
        MZB_CUDA(cudaGetLastError());
        return MZB_OK;
    }

    int initial(const float* d_obs, int B, const int* d_out_slot, const NetOutputs& out, cudaStream_t s) override {
        MZB_REQUIRE(uploaded_, "weights not loaded (mzb_load_weights)");
        MZB_REQUIRE(B <= max_batch_, "batch %d above engine capacity %d", B, max_batch_);
        const NetDims& d = plan_dims_;
        launches_ = 0;
        // NCHW obs -> NHWC bufA
        long total = (long)B * d.initial_flat();
        nchw_to_nhwc_kernel<T><<<(unsigned)((total + 255) / 256), 256, 0, s>>>(d_obs, bufA_.as<T>(), nullptr, (long)d.initial_flat(), B, d.cin, d.in_h * d.in_w);
        ++launches_;
        T* cur = bufA_.as<T>();
        int rc = run_tower(0, B, cur, nullptr, 0, nullptr, s, &cur);
        if (rc) return rc;
        // normalise into the pool slots, then prediction from the pool
        normalize_kernel<T><<<B, 128, 0, s>>>(cur, pool_.as<T>(), d_out_slot, (long)slot_elems_, d.hw(), d.C, normalize_ ? 1 : 0);
        ++launches_;
        rc = run_prediction(B, d_out_slot, out, s);
        if (rc) return rc;
        initial_reward_kernel<<<(B + 127) / 128, 128, 0, s>>>(out.reward, B, renc_.support, (float)renc_.step, (float)renc_.min_value, renc_.contractive);
        ++launches_;
        MZB_CUDA(cudaGetLastError());
        launches_initial_ = launches_;
        return MZB_OK;
    }

    int recurrent(int B, const int* d_in_slot, const int* d_action, const int* d_out_slot, const NetOutputs& out, cudaStream_t s) override {
        MZB_REQUIRE(uploaded_, "weights not loaded (mzb_load_weights)");
        MZB_REQUIRE(B <= max_batch_, "batch %d above engine capacity %d", B, max_batch_);
        const NetDims& d = plan_dims_;
        launches_ = 0;
        T* cur = nullptr;
        int rc = run_tower(1, B, pool_.as<T>(), d_in_slot, (long)slot_elems_, d_action, s, &cur);
        if (rc) return rc;
        // reward head sees the UN-normalised dynamics output (models.cpp:111-116)
        rc = run_head(0, cur, nullptr, (long)slot_elems_, B, out.reward, 1, s);
        if (rc) return rc;
        normalize_kernel<T><<<B, 128, 0, s>>>(cur, pool_.as<T>(), d_out_slot, (long)slot_elems_, d.hw(), d.C, normalize_ ? 1 : 0);
        ++launches_;
        rc = run_prediction(B, d_out_slot, out, s);
        if (rc) return rc;
        MZB_CUDA(cudaGetLastError());
        launches_recurrent_ = launches_;
        return MZB_OK;
    }
    int launches_initial() const override { return launches_initial_; }
    int launches_recurrent() const override { return launches_recurrent_; }

protected:
    // hooks for derived backends
    virtual int pack_extra(const NetPlan&, std::vector<uint8_t>*) { return MZB_OK; }
    // return 1 when the call was executed by a specialised kernel, 0 to fall through to the direct kernel, <0 on error
    virtual int fast_conv(const ConvCall<T>&, cudaStream_t) { return 0; }
    // elements per sample of the activation scratch buffers (a backend may keep activations in a padded layout)
    virtual size_t act_elems(size_t largest_activation) const { return largest_activation; }
    // same contract for AvgPool2d(3, 2, 1) on a compact NHWC activation
    virtual int fast_pool(const T*, T*, int, int, int, int, int, int, cudaStream_t) { return 0; }
    // a run of tower steps starting at step `si` taken by one specialised launch: compact activation in -> compact activation
    // out; *used = steps consumed (0 = not taken)
    virtual int fast_steps(int /*tower*/, size_t /*si*/, int /*B*/, const T* /*in*/, T* /*out*/, cudaStream_t, size_t* used) { *used = 0; return 0; }

    static int conv_smem(int stride) { int PW = TILE * stride + 2; return (CK * PW * PW + 9 * CK * TC) * (int)sizeof(float); }
    int head_smem(int h) const { return (int)head_smem_bytes(heads_[h], head_maxw_[h], head_maxw2_[h]); }

    int launch_conv(const ConvDev& c, int B, const T* in, const int* in_slot, long in_stride, T* out,
                    const T* res, const int* res_slot, long res_stride, const int* action, cudaStream_t s) {
        if (c.has_table && !action) { set_error("conv with an action table needs actions"); return MZB_ERR_STATE; }
        ConvCall<T> call;
        call.conv = &c; call.B = B; call.in = in; call.in_slot = in_slot; call.in_stride = in_stride;
        call.out = out; call.out_slot = nullptr; call.out_stride = (long)c.out_h * c.out_w * c.cout;
        call.res = res; call.res_slot = res_slot; call.res_stride = res ? res_stride : 0;
        call.action = c.has_table ? action : nullptr;
        int fr = fast_conv(call, s);
        if (fr < 0) return fr;
        if (fr == 1) { ++launches_; return MZB_OK; }
        ConvArgs<T> a;
        const float* base = blob_.template as<float>();
        a.in = in; a.in_slot = in_slot; a.in_stride = in_stride;
        a.out = out; a.out_slot = nullptr; a.out_stride = call.out_stride;
        a.res = res; a.res_slot = res_slot; a.res_stride = call.res_stride;
        a.w = base + c.w_off; a.bias = base + c.bias_off;
        a.act_table = c.has_table ? base + c.table_off : nullptr;
        a.action = action;
        a.cin = c.cin; a.cout = c.cout; a.stride = c.stride; a.in_h = c.in_h; a.in_w = c.in_w;
        a.out_h = c.out_h; a.out_w = c.out_w; a.relu = c.relu;
        dim3 grid(((c.out_h + TILE - 1) / TILE) * ((c.out_w + TILE - 1) / TILE), (c.cout + TC - 1) / TC, B);
        conv3_direct_kernel<T><<<grid, 256, conv_smem(c.stride), s>>>(a);
        ++launches_;
        MZB_CUDA(cudaGetLastError());
        return MZB_OK;
    }

    // Runs tower t.  Input: `in` (slot-addressed when in_slot != nullptr).  Scratch ping-pong: bufA/bufB/bufC.
    // On return *cur_out points at the final activation (compact [B][...]).
    // Steps [step_begin, step_end) only when given: `in` is then the activation the step before step_begin produced.
    int run_tower(int t, int B, T* in, const int* in_slot, long in_stride, const int* action, cudaStream_t s, T** cur_out,
                  size_t step_begin = 0, size_t step_end = (size_t)-1) {
        T* bufs[3] = {bufA_.as<T>(), bufB_.as<T>(), bufC_.as<T>()};
        const T* cur = in;
        const int* cur_slot = in_slot;
        long cur_stride = in_stride;
        int cur_c = towers_[t]->in_c, cur_h = towers_[t]->in_h, cur_w = towers_[t]->in_w;
        auto pick = [&](const T* a, const T* b) -> T* {
            for (int i = 0; i < 3; ++i) if (bufs[i] != a && bufs[i] != b) return bufs[i];
            return nullptr;
        };
        for (size_t si = 0; si < steps_[t].size() && si < step_end; ++si) {
            const TowerStep& st = steps_[t][si];
            if (si < step_begin) {   // shape bookkeeping only
                if (st.kind == TowerStep::CONV) { const ConvDev& c = convs_[t][st.conv0]; cur_c = c.cout; cur_h = c.out_h; cur_w = c.out_w; }
                else if (st.kind == TowerStep::POOL) { cur_h = (cur_h + 2 - 3) / 2 + 1; cur_w = (cur_w + 2 - 3) / 2 + 1; }
                continue;
            }
            const long compact = (long)cur_c * cur_h * cur_w;
            const long istride = cur_slot ? cur_stride : compact;
            if (!cur_slot) {
                size_t used = 0;
                T* o = pick(cur, nullptr);
                int fr = fast_steps(t, si, B, cur, o, s, &used);
                if (fr < 0) return fr;
                if (used > 0) {
                    for (size_t k = si; k < si + used; ++k) {
                        const TowerStep& sk = steps_[t][k];
                        if (sk.kind == TowerStep::CONV) { const ConvDev& c = convs_[t][sk.conv0]; cur_c = c.cout; cur_h = c.out_h; cur_w = c.out_w; }
                        else if (sk.kind == TowerStep::POOL) { cur_h = (cur_h + 2 - 3) / 2 + 1; cur_w = (cur_w + 2 - 3) / 2 + 1; }
                    }
                    ++launches_;
                    cur = o;
                    si += used - 1;
                    continue;
                }
            }
            if (st.kind == TowerStep::CONV) {
                const ConvDev& c = convs_[t][st.conv0];
                T* o = pick(cur, nullptr);
                int rc = launch_conv(c, B, cur, cur_slot, istride, o, nullptr, nullptr, 0, action, s);
                if (rc) return rc;
                cur = o; cur_slot = nullptr; cur_c = c.cout; cur_h = c.out_h; cur_w = c.out_w;
            } else if (st.kind == TowerStep::BLOCK) {
                const ConvDev& c1 = convs_[t][st.conv0];
                const ConvDev& c2 = convs_[t][st.conv1];
                T* mid = pick(cur, nullptr);
                int rc = launch_conv(c1, B, cur, cur_slot, istride, mid, nullptr, nullptr, 0, nullptr, s);
                if (rc) return rc;
                // residual = the block input, read where it lies (pool slots or a compact buffer)
                T* o = pick(cur, mid);
                rc = launch_conv(c2, B, mid, nullptr, compact, o, cur, cur_slot, istride, nullptr, s);
                if (rc) return rc;
                cur = o; cur_slot = nullptr;
            } else {  // POOL
                if (cur_slot) { set_error("pooling a slot-addressed activation is not supported"); return MZB_ERR_STATE; }
                int oh = (cur_h + 2 - 3) / 2 + 1, ow = (cur_w + 2 - 3) / 2 + 1;
                T* o = pick(cur, nullptr);
                int fr = fast_pool(cur, o, B, cur_c, cur_h, cur_w, oh, ow, s);
                if (fr < 0) return fr;
                if (fr == 0) {
                    long total = (long)B * oh * ow * cur_c;
                    avgpool3_kernel<T><<<(unsigned)((total + 255) / 256), 256, 0, s>>>(cur, o, B, cur_c, cur_h, cur_w, oh, ow);
                }
                ++launches_;
                cur = o; cur_h = oh; cur_w = ow;
            }
        }
        if (cur_slot) {  // tower without any step (0 blocks): make the activation compact
            T* xc = pick(cur, nullptr);
            long elems = (long)cur_c * cur_h * cur_w, total = (long)B * elems;
            copy_slots_kernel<T><<<(unsigned)((total + 255) / 256), 256, 0, s>>>(cur, cur_slot, cur_stride, xc, elems, total);
            ++launches_;
            cur = xc;
        }
        MZB_CUDA(cudaGetLastError());
        *cur_out = const_cast<T*>(cur);
        return MZB_OK;
    }

    int run_head(int h, const T* x, const int* slots, long stride, int B, double* out, int out_stride, cudaStream_t s) {
        head_kernel<T><<<(B + HEAD_GS - 1) / HEAD_GS, HEAD_THREADS, head_smem(h), s>>>(x, slots, stride, B, heads_[h], out, out_stride, head_maxw_[h], head_maxw2_[h]);
        ++launches_;
        MZB_CUDA(cudaGetLastError());
        return MZB_OK;
    }
    int run_prediction(int B, const int* d_slot, const NetOutputs& out, cudaStream_t s) {
        T* cur = nullptr;
        int rc = run_tower(2, B, pool_.as<T>(), d_slot, (long)slot_elems_, nullptr, s, &cur);
        if (rc) return rc;
        rc = run_head(1, cur, nullptr, (long)slot_elems_, B, out.policy, plan_dims_.A, s);
        if (rc) return rc;
        return run_head(2, cur, nullptr, (long)slot_elems_, B, out.value, 1, s);
    }

    const NetPlan* plan_ = nullptr;
    NetDims plan_dims_;
    bool normalize_ = true, uploaded_ = false;
    const Tower* towers_[3] = {nullptr, nullptr, nullptr};
    std::vector<TowerStep> steps_[3];
    std::vector<ConvDev> convs_[3];
    int max_act_[3] = {0, 0, 0};
    HeadDev heads_[3];
    HeadOff hoff_[3];
    int head_maxw_[3] = {0, 0, 0}, head_maxw2_[3] = {0, 0, 0};
    EncoderConsts renc_;
    DevBuf blob_, pool_, bufA_, bufB_, bufC_;
    size_t fp32_bytes_ = 0;
    size_t slot_elems_ = 0;
    int max_batch_ = 0;
    int launches_ = 0, launches_initial_ = 0, launches_recurrent_ = 0;
};

}  // namespace direct
}  // namespace mzb
