// learn.cu — the forward half of VPRNetModel::Learn (muzero-cpp/vprnet.cpp:192-311; SURVEY.md 8f rank 2, first slice): the K + 1
// unrolled network calls of a training step in TRAIN mode and MuzeroNetworkImpl::loss (models.cpp:268-276), fp32, on the GPU.
//
//   initial_inference(stacked observations) -> loss against step 0 targets (the reward term of step 0 is zero, vprnet.cpp:255)
//   for step 1 .. K: recurrent_inference(previous state, action[step]) -> loss against step targets
//   value / reward targets through ValueEncoder::encode (util.cpp:59-76), losses = sum(-target * log_softmax(logits)) per sample
//   total = mean(priorities [B,1] * (value_loss * w + reward_loss + policy_loss) [B])  — a [B,B] broadcast in the reference, i.e.
//   mean(priorities) * mean(per-sample loss); kept as it is (vprnet.cpp:295)
//   errors = decode(step-0 value logits) - target_values[:, 0]   (new priorities for the replay buffer, vprnet.cpp:259-265)
//
// TRAIN mode means BatchNorm normalises with the statistics of the batch in hand (biased variance, eps 1e-4) and moves the running
// statistics by momentum 0.01 with the unbiased variance (model_layers.cpp:36-42) — every application separately, so the
// dynamics / prediction buffers move K / K + 1 times per call, in the reference's order.  Nothing is folded: the kernels below take
// the module's own tensors (the ones mzb_load_weights / mzb_load_checkpoint received), NCHW fp32 like libtorch, one plain kernel
// per op with fp64 accumulation.  This is the check-mode slice the backward pass and Adam will be verified against; it is not a
// throughput path.  The gradient hooks of Learn (0.5 on the hidden state, gradient_scale on the step losses) only act in backward.
#include <cmath>
#include <map>
#include <string>

#include "engine.h"

namespace mzb {
namespace {

struct T4 { float* p = nullptr; int n = 0, c = 0, h = 0, w = 0; long numel() const { return (long)n * c * h * w; } };

// ---------------------------------------------------------------- kernels (one thread per output element)
__global__ void k_conv3x3(const float* __restrict__ x, const float* __restrict__ w, float* __restrict__ y, int N, int Cin, int H, int W, int Cout,
                          int Ho, int Wo, int stride) {
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long)N * Cout * Ho * Wo) return;
    const int xo = (int)(i % Wo), yo = (int)((i / Wo) % Ho), co = (int)((i / ((long)Wo * Ho)) % Cout), n = (int)(i / ((long)Wo * Ho * Cout));
    double acc = 0.0;
    for (int ci = 0; ci < Cin; ++ci) {
        const float* xp = x + ((long)n * Cin + ci) * H * W;
        const float* wp = w + ((long)co * Cin + ci) * 9;
        for (int ky = 0; ky < 3; ++ky) {
            const int yy = yo * stride + ky - 1;
            if (yy < 0 || yy >= H) continue;
            for (int kx = 0; kx < 3; ++kx) {
                const int xx = xo * stride + kx - 1;
                if (xx < 0 || xx >= W) continue;
                acc += (double)xp[yy * W + xx] * (double)wp[ky * 3 + kx];
            }
        }
    }
    y[i] = (float)acc;
}
__global__ void k_conv1x1(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ b, float* __restrict__ y, int N, int Cin,
                          int HW, int Cout) {
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long)N * Cout * HW) return;
    const int p = (int)(i % HW), co = (int)((i / HW) % Cout), n = (int)(i / ((long)HW * Cout));
    double acc = (double)b[co];
    for (int ci = 0; ci < Cin; ++ci) acc += (double)x[((long)n * Cin + ci) * HW + p] * (double)w[(long)co * Cin + ci];
    y[i] = (float)acc;
}
// BatchNorm2d in training mode: one block per channel.  stats over N*H*W in fp64; y = (x - mean) * rsqrt(var_biased + eps) * gamma + beta
// (+ residual) (ReLU); running statistics moved by `momentum` with the unbiased variance.
__global__ void k_bn_train(const float* __restrict__ x, float* __restrict__ y, const float* __restrict__ gamma, const float* __restrict__ beta,
                           float* running_mean, float* running_var, const float* __restrict__ residual, int N, int C, int HW, float eps,
                           float momentum, int relu) {
    const int c = blockIdx.x;
    __shared__ double s_sum[256], s_sq[256];
    const long cnt = (long)N * HW;
    double sum = 0.0, sq = 0.0;
    for (long i = threadIdx.x; i < cnt; i += blockDim.x) {
        const double v = (double)x[((i / HW) * C + c) * HW + (i % HW)];
        sum += v; sq += v * v;
    }
    s_sum[threadIdx.x] = sum; s_sq[threadIdx.x] = sq;
    __syncthreads();
    for (int o = blockDim.x / 2; o > 0; o >>= 1) {
        if ((int)threadIdx.x < o) { s_sum[threadIdx.x] += s_sum[threadIdx.x + o]; s_sq[threadIdx.x] += s_sq[threadIdx.x + o]; }
        __syncthreads();
    }
    const double mean = s_sum[0] / (double)cnt;
    double var = s_sq[0] / (double)cnt - mean * mean;
    if (var < 0.0) var = 0.0;
    const float fmean = (float)mean, inv = (float)(1.0 / sqrt(var + (double)eps));
    const float g = gamma[c], b = beta[c];
    for (long i = threadIdx.x; i < cnt; i += blockDim.x) {
        const long at = ((i / HW) * C + c) * HW + (i % HW);
        float v = (x[at] - fmean) * inv * g + b;
        if (residual) v += residual[at];
        if (relu) v = fmaxf(v, 0.f);
        y[at] = v;
    }
    if (threadIdx.x == 0) {
        const double unbiased = cnt > 1 ? var * (double)cnt / (double)(cnt - 1) : var;
        running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * fmean;
        running_var[c] = (1.f - momentum) * running_var[c] + momentum * (float)unbiased;
    }
}
__global__ void k_linear(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ b, float* __restrict__ y, int N, int In, int Out, int relu) {
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long)N * Out) return;
    const int o = (int)(i % Out), n = (int)(i / Out);
    double acc = (double)b[o];
    for (int k = 0; k < In; ++k) acc += (double)x[(long)n * In + k] * (double)w[(long)o * In + k];
    float v = (float)acc;
    if (relu) v = fmaxf(v, 0.f);
    y[i] = v;
}
__global__ void k_avgpool3s2(const float* __restrict__ x, float* __restrict__ y, int NC, int H, int W, int Ho, int Wo) {   // AvgPool2d(3, 2, 1), pad counted
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long)NC * Ho * Wo) return;
    const int xo = (int)(i % Wo), yo = (int)((i / Wo) % Ho);
    const long nc = i / ((long)Wo * Ho);
    float acc = 0.f;
    for (int ky = 0; ky < 3; ++ky)
        for (int kx = 0; kx < 3; ++kx) {
            const int yy = yo * 2 + ky - 1, xx = xo * 2 + kx - 1;
            if (yy >= 0 && yy < H && xx >= 0 && xx < W) acc += x[(nc * H + yy) * W + xx];
        }
    y[i] = acc / 9.f;
}
// normalize_hidden_state (models.cpp:16-26): per (sample, channel) min-max over the pixels
__global__ void k_normalize(const float* __restrict__ x, float* __restrict__ y, long NC, int HW) {
    const long r = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= NC) return;
    const float* row = x + r * HW;
    float mn = row[0], mx = row[0];
    for (int i = 1; i < HW; ++i) { mn = fminf(mn, row[i]); mx = fmaxf(mx, row[i]); }
    float scale = mx - mn;
    if (scale < 1e-5f) scale = 1e-5f;
    for (int i = 0; i < HW; ++i) y[r * HW + i] = (row[i] - mn) / scale;
}
// cat({state, action planes}, 1): planes[action[n, step]] appended as channels C .. C + ach - 1
__global__ void k_cat_action(const float* __restrict__ state, const float* __restrict__ planes, const int* __restrict__ actions, int a_stride, int step,
                             float* __restrict__ y, int N, int C, int ach, int HW) {
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long)N * (C + ach) * HW) return;
    const int p = (int)(i % HW), c = (int)((i / HW) % (C + ach)), n = (int)(i / ((long)HW * (C + ach)));
    y[i] = c < C ? state[((long)n * C + c) * HW + p] : planes[((long)actions[(long)n * a_stride + step] * ach + (c - C)) * HW + p];
}
// loss[n] += sum_k -target[n, k] * log_softmax(logits[n])[k]   (models.cpp:271-273)
__global__ void k_ce_accumulate(const float* __restrict__ logits, const float* __restrict__ target, int t_stride, double* loss, int N, int S) {
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    const float* lg = logits + (long)n * S;
    const float* tg = target + (long)n * t_stride;
    float mx = lg[0];
    for (int k = 1; k < S; ++k) mx = fmaxf(mx, lg[k]);
    float sum = 0.f;
    for (int k = 0; k < S; ++k) sum += expf(lg[k] - mx);
    const float lse = mx + logf(sum);
    float acc = 0.f;
    for (int k = 0; k < S; ++k) acc += -tg[k] * (lg[k] - lse);
    loss[n] += (double)acc;
}

inline unsigned blocks(long n) { return (unsigned)((n + 255) / 256); }

// ---------------------------------------------------------------- the train-mode network over the module's own tensors
struct LearnNet {
    mzb_engine* e;
    cudaStream_t s;
    std::map<std::string, DevBuf> W;        // name -> device copy (BatchNorm buffers are updated in place)
    std::vector<std::unique_ptr<DevBuf>> scratch;
    std::string err;

    float* w(const std::string& name) {
        auto it = W.find(name);
        if (it == W.end()) { if (err.empty()) err = "missing tensor " + name; return nullptr; }
        return it->second.as<float>();
    }
    T4 alloc(int n, int c, int h, int ww) {
        T4 t; t.n = n; t.c = c; t.h = h; t.w = ww;
        scratch.emplace_back(new DevBuf());
        if (scratch.back()->alloc((size_t)t.numel() * sizeof(float)) != cudaSuccess) { if (err.empty()) err = "out of device memory"; return t; }
        t.p = scratch.back()->as<float>();
        return t;
    }
    T4 conv3(const T4& x, const std::string& name, int cout, int stride) {
        T4 y = alloc(x.n, cout, (x.h + 2 - 3) / stride + 1, (x.w + 2 - 3) / stride + 1);
        const float* wt = w(name + ".weight");
        if (wt && y.p) k_conv3x3<<<blocks(y.numel()), 256, 0, s>>>(x.p, wt, y.p, x.n, x.c, x.h, x.w, cout, y.h, y.w, stride);
        return y;
    }
    T4 bn(const T4& x, const std::string& name, const T4* residual, bool relu) {
        T4 y = alloc(x.n, x.c, x.h, x.w);
        float *g = w(name + ".weight"), *b = w(name + ".bias"), *rm = w(name + ".running_mean"), *rv = w(name + ".running_var");
        if (g && b && rm && rv && y.p)
            k_bn_train<<<x.c, 256, 0, s>>>(x.p, y.p, g, b, rm, rv, residual ? residual->p : nullptr, x.n, x.c, x.h * x.w, 1e-4f, 0.01f, relu ? 1 : 0);
        return y;
    }
    // ResidualBlock (model_layers.cpp:70-88): relu(bn2(conv2(relu(bn1(conv1 x)))) + x)
    T4 block(const T4& x, const std::string& prefix, int idx) {
        const std::string p = prefix + ".resnet_" + std::to_string(idx) + "_";
        T4 a = bn(conv3(x, p + "conv1", x.c, 1), p + "bn1", nullptr, true);
        return bn(conv3(a, p + "conv2", x.c, 1), p + "bn2", &x, true);
    }
    T4 normalize(const T4& x) {
        if (!e->cfg.normalize_hidden_states) return x;
        T4 y = alloc(x.n, x.c, x.h, x.w);
        if (y.p) k_normalize<<<blocks((long)x.n * x.c), 256, 0, s>>>(x.p, y.p, (long)x.n * x.c, x.h * x.w);
        return y;
    }
    // conv1x1 -> flatten (channel-major, as .view() of an NCHW tensor) -> MLP (model_layers.cpp:46-65)
    T4 head(const T4& x, const std::string& conv, const std::string& mlp, int red, const int32_t* layers, int n_layers, int out) {
        T4 r = alloc(x.n, red, x.h, x.w);
        float *cw = w(conv + ".weight"), *cb = w(conv + ".bias");
        if (cw && cb && r.p) k_conv1x1<<<blocks(r.numel()), 256, 0, s>>>(x.p, cw, cb, r.p, x.n, x.c, x.h * x.w, red);
        T4 cur = r;
        int in = red * x.h * x.w;
        for (int l = 0; l <= n_layers; ++l) {
            const int o = l < n_layers ? layers[l] : out;
            T4 y = alloc(x.n, o, 1, 1);
            float *lw = w(mlp + ".linear_" + std::to_string(l) + ".weight"), *lb = w(mlp + ".linear_" + std::to_string(l) + ".bias");
            if (lw && lb && y.p) k_linear<<<blocks((long)x.n * o), 256, 0, s>>>(cur.p, lw, lb, y.p, x.n, in, o, l < n_layers ? 1 : 0);
            cur = y; in = o;
        }
        return cur;
    }
    T4 pool(const T4& x) {
        T4 y = alloc(x.n, x.c, (x.h + 2 - 3) / 2 + 1, (x.w + 2 - 3) / 2 + 1);
        if (y.p) k_avgpool3s2<<<blocks(y.numel()), 256, 0, s>>>(x.p, y.p, x.n * x.c, x.h, x.w, y.h, y.w);
        return y;
    }
    // RepresentationNetworkImpl::forward (models.cpp:31-65) + normalize (models.cpp:227-229)
    T4 representation(const T4& obs) {
        const mzb_config& c = e->cfg;
        const int C = c.resnet_channels;
        const std::string rh = "representation_network.representation_head.0";
        T4 x;
        if (c.downsample) {   // model_layers.cpp:140-155
            x = conv3(obs, rh + ".resnet_head_downsample_conv1", C / 2, 2);
            for (int i = 0; i < 2; ++i) x = block(x, rh + ".resnet_head_downsample_resblock1." + std::to_string(i), i);
            x = conv3(x, rh + ".resnet_head_downsample_conv2", C, 2);
            for (int i = 2; i < 5; ++i) x = block(x, rh + ".resnet_head_downsample_resblock2." + std::to_string(i - 2), i);
            x = pool(x);
            for (int i = 5; i < 8; ++i) x = block(x, rh + ".resnet_head_downsample_resblock3." + std::to_string(i - 5), i);
            x = pool(x);
        } else {
            x = bn(conv3(obs, rh + ".representation_resnet_head_conv", C, 1), rh + ".representation_resnet_head_bn", nullptr, true);
        }
        for (int i = 0; i < c.representation_blocks; ++i) x = block(x, "representation_network.representation_layers." + std::to_string(i), i);
        return normalize(x);
    }
    // PredictionNetworkImpl::forward (models.cpp:153-166)
    void prediction(const T4& state, T4* policy, T4* value) {
        const mzb_config& c = e->cfg;
        T4 x = state;
        for (int i = 0; i < c.prediction_blocks; ++i) x = block(x, "prediction_network.prediction_resnet_body." + std::to_string(i), i);
        *policy = head(x, "prediction_network.prediction_policy_conv", "prediction_network.prediction_policy_mlp.prediction_policy_mlp",
                       c.policy_reduced_channels, c.policy_head_layers, c.n_policy_head_layers, e->dims.A);
        *value = head(x, "prediction_network.prediction_value_conv", "prediction_network.prediction_value_mlp.prediction_value_mlp",
                      c.value_reduced_channels, c.value_head_layers, c.n_value_head_layers, e->dims.Sv);
    }
    // DynamicsNetworkImpl::forward (models.cpp:105-117): reward from the UN-normalised output, next state normalised (models.cpp:236-238)
    T4 dynamics(const T4& state, const float* planes, const int* actions, int a_stride, int step, T4* reward) {
        const mzb_config& c = e->cfg;
        const int C = c.resnet_channels, ach = c.action_channels;
        T4 in = alloc(state.n, C + ach, state.h, state.w);
        if (in.p) k_cat_action<<<blocks(in.numel()), 256, 0, s>>>(state.p, planes, actions, a_stride, step, in.p, state.n, C, ach, state.h * state.w);
        const std::string dh = "dynamics_network.dynamics_resnet_head";
        T4 x = bn(conv3(in, dh + ".resnet_head_conv", C, 1), dh + ".resnet_head_bn", nullptr, true);
        for (int i = 0; i < c.dynamics_blocks; ++i) x = block(x, "dynamics_network.dynamics_resnet_body." + std::to_string(i), i);
        *reward = head(x, "dynamics_network.dynamics_reward_conv", "dynamics_network.dynamics_reward_mlp.dynamics_reward_mlp",
                       c.reward_reduced_channels, c.reward_head_layers, c.n_reward_head_layers, e->dims.Sr);
        return normalize(x);
    }
};

}  // namespace
}  // namespace mzb

using namespace mzb;

extern "C" {

int mzb_learn_forward(mzb_engine* e, const mzb_batch* batch, int num_unroll_steps, double value_loss_weight, mzb_loss_info* out) {
    MZB_REQUIRE(e && batch && out, "null argument");
    MZB_REQUIRE(e->cfg.evaluator == MZB_EVAL_NETWORK && e->weights_loaded && !e->raw_weights.empty(), "no weights loaded (mzb_load_weights / mzb_load_checkpoint)");
    const int B = batch->num_samples, K1 = num_unroll_steps + 1, A = e->dims.A;
    MZB_REQUIRE(B >= 1 && num_unroll_steps >= 0, "empty batch or negative unroll");
    MZB_REQUIRE(batch->actions && batch->stacked_observations && batch->target_values && batch->target_rewards && batch->target_policies && batch->priorities,
                "null batch field");
    for (int i = 0; i < B * K1; ++i) MZB_REQUIRE(batch->actions[i] >= 0 && batch->actions[i] < A, "batch action %d outside the action space", batch->actions[i]);
    MZB_CUDA(cudaSetDevice(e->device));
    {   // tensors trained by mzb_learn_step since the last look are the current ones
        const int src = learn_sync_host(e);
        if (src) return src;
    }
    try {
        LearnNet net;
        net.e = e; net.s = e->stream;
        // the module's own tensors; BatchNorm running statistics persist across calls in e->learn_state (a training run moves them
        // step by step) — seeded from the loaded tensors on the first call after a load
        if (e->learn_state.empty() || e->learn_state_serial != e->weights_serial) {
            e->learn_state.clear();
            for (const mzb_engine::RawTensor& r : e->raw_weights) e->learn_state[r.name] = r.data;
            e->learn_state_serial = e->weights_serial;
        }
        for (auto& kv : e->learn_state) {
            DevBuf& d = net.W[kv.first];
            MZB_CUDA(d.alloc(std::max<size_t>(kv.second.size(), 1) * sizeof(float)));
            MZB_CUDA(cudaMemcpyAsync(d.p, kv.second.data(), kv.second.size() * sizeof(float), cudaMemcpyHostToDevice, net.s));
        }
        const NetDims& d = e->dims;
        // inputs
        T4 obs = net.alloc(B, d.cin, d.in_h, d.in_w);
        MZB_REQUIRE(obs.p, "out of device memory");
        MZB_CUDA(cudaMemcpyAsync(obs.p, batch->stacked_observations, (size_t)obs.numel() * sizeof(float), cudaMemcpyHostToDevice, net.s));
        DevBuf d_actions, d_planes, d_tv, d_tr, d_tp, d_loss;
        MZB_CUDA(d_actions.alloc((size_t)B * K1 * sizeof(int)));
        MZB_CUDA(cudaMemcpyAsync(d_actions.p, batch->actions, (size_t)B * K1 * sizeof(int), cudaMemcpyHostToDevice, net.s));
        // action_representation_recurrent per action (vprnet.cpp:201-206): the engine's planes, or the examples' constant plane a / A
        const int hw = d.hw();
        std::vector<float> planes((size_t)A * std::max(1, d.ach) * hw);
        for (int a = 0; a < A; ++a)
            for (int i = 0; i < d.ach * hw; ++i)
                planes[(size_t)a * d.ach * hw + i] = e->cfg.action_planes ? e->cfg.action_planes[(size_t)a * d.ach * hw + i] : (float)((double)a / A);
        MZB_CUDA(d_planes.alloc(planes.size() * sizeof(float)));
        MZB_CUDA(cudaMemcpyAsync(d_planes.p, planes.data(), planes.size() * sizeof(float), cudaMemcpyHostToDevice, net.s));
        // targets: scalars -> float -> categorical support (ValueEncoder::encode), policies -> float
        const int Sv = d.Sv, Sr = d.Sr;
        std::vector<float> tv((size_t)B * K1), tr((size_t)B * K1), tp((size_t)B * K1 * A), tv_enc((size_t)B * K1 * Sv), tr_enc((size_t)B * K1 * Sr);
        for (size_t i = 0; i < tv.size(); ++i) { tv[i] = (float)batch->target_values[i]; tr[i] = (float)batch->target_rewards[i]; }
        for (size_t i = 0; i < tp.size(); ++i) tp[i] = (float)batch->target_policies[i];
        MZB_CUDA(cudaStreamSynchronize(net.s));
        int rc = mzb_value_encode(e, 0, tv.data(), B * K1, tv_enc.data());
        if (rc) return rc;
        rc = mzb_value_encode(e, 1, tr.data(), B * K1, tr_enc.data());
        if (rc) return rc;
        MZB_CUDA(d_tv.alloc(tv_enc.size() * sizeof(float)));
        MZB_CUDA(d_tr.alloc(tr_enc.size() * sizeof(float)));
        MZB_CUDA(d_tp.alloc(tp.size() * sizeof(float)));
        MZB_CUDA(cudaMemcpyAsync(d_tv.p, tv_enc.data(), tv_enc.size() * sizeof(float), cudaMemcpyHostToDevice, net.s));
        MZB_CUDA(cudaMemcpyAsync(d_tr.p, tr_enc.data(), tr_enc.size() * sizeof(float), cudaMemcpyHostToDevice, net.s));
        MZB_CUDA(cudaMemcpyAsync(d_tp.p, tp.data(), tp.size() * sizeof(float), cudaMemcpyHostToDevice, net.s));
        MZB_CUDA(d_loss.alloc((size_t)3 * B * sizeof(double)));   // value | reward | policy, per sample
        MZB_CUDA(cudaMemsetAsync(d_loss.p, 0, (size_t)3 * B * sizeof(double), net.s));
        double* lv = d_loss.as<double>(); double* lr = lv + B; double* lp = lr + B;
        const unsigned gb = (unsigned)((B + 127) / 128);

        // step 0: initial inference (models.cpp:222-229); its reward term is zeros_like (vprnet.cpp:255)
        T4 state = net.representation(obs), policy, value, reward;
        net.prediction(state, &policy, &value);
        MZB_REQUIRE(net.err.empty(), "%s", net.err.c_str());
        k_ce_accumulate<<<gb, 128, 0, net.s>>>(value.p, d_tv.as<float>(), K1 * Sv, lv, B, Sv);
        k_ce_accumulate<<<gb, 128, 0, net.s>>>(policy.p, d_tp.as<float>(), K1 * A, lp, B, A);
        std::vector<float> v0((size_t)B * Sv);
        MZB_CUDA(cudaMemcpyAsync(v0.data(), value.p, v0.size() * sizeof(float), cudaMemcpyDeviceToHost, net.s));
        // steps 1 .. K: recurrent inference from the previous state (vprnet.cpp:268-291)
        for (int step = 1; step < K1; ++step) {
            state = net.dynamics(state, d_planes.as<float>(), d_actions.as<int>(), K1, step, &reward);
            net.prediction(state, &policy, &value);
            MZB_REQUIRE(net.err.empty(), "%s", net.err.c_str());
            k_ce_accumulate<<<gb, 128, 0, net.s>>>(value.p, d_tv.as<float>() + (size_t)step * Sv, K1 * Sv, lv, B, Sv);
            k_ce_accumulate<<<gb, 128, 0, net.s>>>(reward.p, d_tr.as<float>() + (size_t)step * Sr, K1 * Sr, lr, B, Sr);
            k_ce_accumulate<<<gb, 128, 0, net.s>>>(policy.p, d_tp.as<float>() + (size_t)step * A, K1 * A, lp, B, A);
        }
        MZB_CUDA(cudaGetLastError());
        std::vector<double> loss((size_t)3 * B);
        MZB_CUDA(cudaMemcpyAsync(loss.data(), d_loss.p, loss.size() * sizeof(double), cudaMemcpyDeviceToHost, net.s));
        // BatchNorm running statistics back into the persistent state
        for (auto& kv : e->learn_state) {
            const std::string& n = kv.first;
            const bool rs = (n.size() > 12 && n.compare(n.size() - 12, 12, "running_mean") == 0) || (n.size() > 11 && n.compare(n.size() - 11, 11, "running_var") == 0);
            if (rs) MZB_CUDA(cudaMemcpyAsync(kv.second.data(), net.W[n].p, kv.second.size() * sizeof(float), cudaMemcpyDeviceToHost, net.s));
        }
        MZB_CUDA(cudaStreamSynchronize(net.s));
        e->learn_host_dirty = true;
        // per-sample sums -> LossInfo (vprnet.cpp:293-310)
        double sv = 0, sr = 0, sp = 0, st = 0, spri = 0;
        for (int b = 0; b < B; ++b) {
            sv += loss[b]; sr += loss[B + b]; sp += loss[2 * B + b];
            st += loss[b] * value_loss_weight + loss[B + b] + loss[2 * B + b];
            spri += batch->priorities[b];
        }
        out->value = sv / B; out->reward = sr / B; out->policy = sp / B;
        out->total_loss = (spri / B) * (st / B);   // mean over the [B,1] x [B] broadcast
        if (out->errors) {
            std::vector<float> dec(B);
            rc = mzb_value_decode(e, 0, v0.data(), B, 1, dec.data());
            if (rc) return rc;
            for (int b = 0; b < B; ++b) out->errors[b] = (double)dec[b] - (double)tv[(size_t)b * K1];
        }
        return MZB_OK;
    } catch (const std::exception& ex) {
        set_error("mzb_learn_forward: %s", ex.what());
        return MZB_ERR_INVALID;
    }
}

// a tensor of the training state (BatchNorm running statistics after mzb_learn_forward calls; everything else as loaded)
int mzb_learn_get_tensor(mzb_engine* e, const char* name, float* out, int64_t numel) {
    MZB_REQUIRE(e && name && out, "null argument");
    {
        const int src = learn_sync_host(e);
        if (src) return src;
    }
    auto it = e->learn_state.find(name);
    MZB_REQUIRE(it != e->learn_state.end(), "no tensor %s in the training state (call mzb_learn_forward first)", name);
    MZB_REQUIRE((int64_t)it->second.size() == numel, "tensor %s has %zu elements, not %lld", name, it->second.size(), (long long)numel);
    memcpy(out, it->second.data(), it->second.size() * sizeof(float));
    return MZB_OK;
}

}  // extern "C"
