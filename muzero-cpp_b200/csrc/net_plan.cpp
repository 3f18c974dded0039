// net_plan.cpp — fold the reference's named module tensors into per-layer kernels' weights (host, fp64 folding).
#include "net_plan.h"

#include <cmath>
#include <cstring>

namespace mzb {

static thread_local std::string g_error;
void set_error(const char* fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    g_error = buf;
}
const char* get_error() { return g_error.c_str(); }

// ValueEncoder::ValueEncoder (util.cpp:45-54).  The reference maps the bounds through h() on a float32 tensor
// (torch::tensor({double}) has the default dtype) and widens the result, so h() is evaluated in float here.
static double contractive_f32(double xd) {
    float x = (float)xd;
    float s = (x > 0.f) ? 1.f : ((x < 0.f) ? -1.f : 0.f);
    float a = std::fabs(x) + 1.0f;
    float r = std::sqrt(a) - 1.0f;
    float left = s * r;
    float right = 0.001f * x;
    return (double)(left + right);
}
EncoderConsts make_encoder(double min_value, double max_value, bool contractive) {
    EncoderConsts e;
    e.contractive = contractive ? 1 : 0;
    if (contractive) {
        max_value = contractive_f32(max_value);
        min_value = contractive_f32(min_value);
    }
    e.min_value = min_value;
    e.max_value = max_value;
    e.support = (int)(std::ceil(max_value) - std::floor(min_value) + 1);
    e.step = (max_value - min_value) / (e.support - 1);
    return e;
}

int derive_dims(const mzb_config& c, NetDims* d, EncoderConsts* venc, EncoderConsts* renc) {
    MZB_REQUIRE(c.num_actions >= 1 && c.num_actions <= MZB_MAX_ACTIONS, "num_actions must be in [1,%d]", MZB_MAX_ACTIONS);
    MZB_REQUIRE(c.num_players == 1 || c.num_players == 2, "only 1 or 2 player games (mcts.cpp:30)");
    MZB_REQUIRE(c.observation_c > 0 && c.observation_h > 0 && c.observation_w > 0, "bad observation_shape");
    MZB_REQUIRE(c.resnet_channels > 0 && c.action_channels >= 0, "bad channel counts");
    // models.cpp:183-187
    d->cin = c.observation_c * (c.stacked_observations + 1) + c.stacked_observations * c.action_channels;
    d->in_h = c.observation_h;
    d->in_w = c.observation_w;
    d->C = c.resnet_channels;
    if (c.downsample) {  // model_layers.cpp:157-162
        d->h = (c.observation_h + 15) / 16;
        d->w = (c.observation_w + 15) / 16;
    } else {
        d->h = c.observation_h;
        d->w = c.observation_w;
    }
    d->A = c.num_actions;
    d->ach = c.action_channels;
    *venc = make_encoder(c.min_value, c.max_value, c.use_contractive_mapping != 0);
    *renc = make_encoder(c.min_reward, c.max_reward, c.use_contractive_mapping != 0);
    d->Sv = venc->support;
    d->Sr = renc->support;
    MZB_REQUIRE(d->Sv <= 64 && d->Sr <= 64, "support size above 64 is not supported");
    return MZB_OK;
}

namespace {
struct Getter {
    const TensorMap& m;
    bool ok = true;
    std::string missing;
    const float* get(const std::string& name, std::initializer_list<int64_t> dims) {
        auto it = m.find(name);
        if (it == m.end()) { if (ok) missing = "missing tensor " + name; ok = false; return nullptr; }
        const auto& have = it->second.second;
        std::vector<int64_t> want(dims);
        if (have != want) { if (ok) missing = "bad shape for " + name; ok = false; return nullptr; }
        return it->second.first;
    }
};

// conv weight [cout][cin][3][3] (+ optional BN) -> Conv3Host.  cin_main < cin means the trailing channels are
// action planes handled through act_table.
void fold_conv(Getter& g, const std::string& conv_name, const std::string& bn_name, int cin_total, int cin_main,
               int cout, int stride, int in_h, int in_w, bool relu, Conv3Host* out) {
    out->cin = cin_main; out->cout = cout; out->stride = stride; out->in_h = in_h; out->in_w = in_w;
    out->out_h = (in_h + 2 - 3) / stride + 1;
    out->out_w = (in_w + 2 - 3) / stride + 1;
    out->relu = relu;
    const float* w = g.get(conv_name + ".weight", {cout, cin_total, 3, 3});
    std::vector<double> scale(cout, 1.0), shift(cout, 0.0);
    if (!bn_name.empty()) {
        const float* gamma = g.get(bn_name + ".weight", {cout});
        const float* beta = g.get(bn_name + ".bias", {cout});
        const float* mean = g.get(bn_name + ".running_mean", {cout});
        const float* var = g.get(bn_name + ".running_var", {cout});
        if (g.ok) {
            for (int o = 0; o < cout; ++o) {
                double s = (double)gamma[o] / std::sqrt((double)var[o] + 1e-4);  // eps: model_layers.cpp:38
                scale[o] = s;
                shift[o] = (double)beta[o] - (double)mean[o] * s;
            }
        }
    }
    if (!g.ok) return;
    out->w.assign((size_t)cout * 9 * cin_main, 0.f);
    out->bias.assign(cout, 0.f);
    for (int o = 0; o < cout; ++o) {
        out->bias[o] = (float)shift[o];
        for (int ci = 0; ci < cin_main; ++ci)
            for (int t = 0; t < 9; ++t)
                out->w[((size_t)o * 9 + t) * cin_main + ci] = (float)((double)w[((size_t)o * cin_total + ci) * 9 + t] * scale[o]);
    }
    // keep raw pointers to the action-plane weights for the table (filled by caller)
}

void add_block(Getter& g, Tower* t, const std::string& prefix, int idx, int C, int h, int w) {
    std::string p = prefix + ".resnet_" + std::to_string(idx) + "_";
    Conv3Host c1, c2;
    fold_conv(g, p + "conv1", p + "bn1", C, C, C, 1, h, w, true, &c1);
    fold_conv(g, p + "conv2", p + "bn2", C, C, C, 1, h, w, true, &c2);  // relu after the residual add
    TowerStep s;
    s.kind = TowerStep::BLOCK;
    s.conv0 = (int)t->convs.size();
    t->convs.push_back(std::move(c1));
    s.conv1 = (int)t->convs.size();
    t->convs.push_back(std::move(c2));
    t->steps.push_back(s);
}
void add_conv(Tower* t, Conv3Host&& c) {
    TowerStep s;
    s.kind = TowerStep::CONV;
    s.conv0 = (int)t->convs.size();
    t->convs.push_back(std::move(c));
    t->steps.push_back(s);
}
void finish_tower(Tower* t, int in_c, int in_h, int in_w) {
    t->in_c = in_c; t->in_h = in_h; t->in_w = in_w;
    int c = in_c, h = in_h, w = in_w;
    int mx = c * h * w;
    for (auto& s : t->steps) {
        if (s.kind == TowerStep::POOL) { h = (h + 2 - 3) / 2 + 1; w = (w + 2 - 3) / 2 + 1; }
        else { const Conv3Host& cv = t->convs[s.kind == TowerStep::BLOCK ? s.conv1 : s.conv0]; c = cv.cout; h = cv.out_h; w = cv.out_w; }
        if (c * h * w > mx) mx = c * h * w;
    }
    t->out_c = c; t->out_h = h; t->out_w = w; t->max_act_elems = mx;
}
void build_head(Getter& g, const std::string& conv, const std::string& mlp, int C, int red, int hw,
                const int32_t* layers, int n_layers, int out_size, HeadHost* h) {
    h->C = C; h->red = red; h->hw = hw;
    const float* w1 = g.get(conv + ".weight", {red, C, 1, 1});
    const float* b1 = g.get(conv + ".bias", {red});
    h->sizes.clear();
    h->sizes.push_back(red * hw);
    for (int i = 0; i < n_layers; ++i) h->sizes.push_back(layers[i]);
    h->sizes.push_back(out_size);
    if (g.ok) { h->w1.assign(w1, w1 + (size_t)red * C); h->b1.assign(b1, b1 + red); }
    h->W.clear(); h->b.clear();
    for (size_t l = 0; l + 1 < h->sizes.size(); ++l) {
        std::string n = mlp + ".linear_" + std::to_string(l);
        const float* W = g.get(n + ".weight", {h->sizes[l + 1], h->sizes[l]});
        const float* b = g.get(n + ".bias", {h->sizes[l + 1]});
        if (g.ok) {
            h->W.emplace_back(W, W + (size_t)h->sizes[l + 1] * h->sizes[l]);
            h->b.emplace_back(b, b + h->sizes[l + 1]);
        }
    }
}
double conv_flops(const Conv3Host& c, int extra_cin = 0) { return 2.0 * (c.cin + extra_cin) * c.cout * 9.0 * c.out_h * c.out_w; }
double tower_flops(const Tower& t) { double f = 0; for (auto& c : t.convs) f += conv_flops(c); return f; }
double head_flops(const HeadHost& h) {
    double f = 2.0 * h.C * h.red * h.hw;
    for (size_t l = 0; l + 1 < h.sizes.size(); ++l) f += 2.0 * h.sizes[l] * h.sizes[l + 1];
    return f;
}
}  // namespace

int build_plan(const mzb_config& cfg, const TensorMap& tensors, NetPlan* plan) {
    int rc = derive_dims(cfg, &plan->dims, &plan->value_enc, &plan->reward_enc);
    if (rc != MZB_OK) return rc;
    const NetDims& d = plan->dims;
    plan->normalize = cfg.normalize_hidden_states != 0;
    Getter g{tensors};
    const int C = d.C;

    // ---- representation (models.cpp:31-65) ----
    Tower& rep = plan->representation;
    rep = Tower();
    const std::string rh = "representation_network.representation_head.0";
    if (cfg.downsample) {  // model_layers.cpp:116-155: conv s2 (no BN/ReLU) -> 2 blocks -> conv s2 -> 3 blocks -> pool -> 3 blocks -> pool
        MZB_REQUIRE(C % 2 == 0, "downsample head needs an even channel count");
        Conv3Host c1;
        fold_conv(g, rh + ".resnet_head_downsample_conv1", "", d.cin, d.cin, C / 2, 2, d.in_h, d.in_w, false, &c1);
        int h1 = c1.out_h, w1 = c1.out_w;
        add_conv(&rep, std::move(c1));
        for (int i = 0; i < 2; ++i) add_block(g, &rep, rh + ".resnet_head_downsample_resblock1." + std::to_string(i), i, C / 2, h1, w1);
        Conv3Host c2;
        fold_conv(g, rh + ".resnet_head_downsample_conv2", "", C / 2, C / 2, C, 2, h1, w1, false, &c2);
        int h2 = c2.out_h, w2 = c2.out_w;
        add_conv(&rep, std::move(c2));
        for (int i = 2; i < 5; ++i) add_block(g, &rep, rh + ".resnet_head_downsample_resblock2." + std::to_string(i - 2), i, C, h2, w2);
        TowerStep pool; pool.kind = TowerStep::POOL;
        rep.steps.push_back(pool);
        int h3 = (h2 + 2 - 3) / 2 + 1, w3 = (w2 + 2 - 3) / 2 + 1;
        for (int i = 5; i < 8; ++i) add_block(g, &rep, rh + ".resnet_head_downsample_resblock3." + std::to_string(i - 5), i, C, h3, w3);
        rep.steps.push_back(pool);
        int h4 = (h3 + 2 - 3) / 2 + 1, w4 = (w3 + 2 - 3) / 2 + 1;
        MZB_REQUIRE(h4 == d.h && w4 == d.w, "downsample head output %dx%d differs from encoded_state_shape %dx%d (the reference has the same mismatch for sizes that are not multiples of 16)", h4, w4, d.h, d.w);
    } else {
        Conv3Host c;
        fold_conv(g, rh + ".representation_resnet_head_conv", rh + ".representation_resnet_head_bn", d.cin, d.cin, C, 1, d.in_h, d.in_w, true, &c);
        add_conv(&rep, std::move(c));
    }
    for (int i = 0; i < cfg.representation_blocks; ++i)
        add_block(g, &rep, "representation_network.representation_layers." + std::to_string(i), i, C, d.h, d.w);
    finish_tower(&rep, d.cin, d.in_h, d.in_w);

    // ---- dynamics (models.cpp:80-117) ----
    Tower& dyn = plan->dynamics;
    dyn = Tower();
    {
        const std::string dh = "dynamics_network.dynamics_resnet_head";
        Conv3Host c;
        fold_conv(g, dh + ".resnet_head_conv", dh + ".resnet_head_bn", C + d.ach, C, C, 1, d.h, d.w, true, &c);
        // action planes: torch::cat({state, action}, 1) (models.cpp:232) never materialised; the planes are
        // constants per action, so their convolution is a per-(action, pixel, cout) additive table.
        if (g.ok && d.ach > 0) {
            const float* w = g.get(dh + ".resnet_head_conv.weight", {C, C + d.ach, 3, 3});
            const float* gamma = g.get(dh + ".resnet_head_bn.weight", {C});
            const float* var = g.get(dh + ".resnet_head_bn.running_var", {C});
            const int hw = d.hw();
            std::vector<float> default_planes;
            const float* planes = cfg.action_planes;
            if (!planes) {  // examples' encode_action: plane filled with (double)action / num_actions
                default_planes.resize((size_t)d.A * d.ach * hw);
                for (int a = 0; a < d.A; ++a)
                    for (int i = 0; i < d.ach * hw; ++i) default_planes[(size_t)a * d.ach * hw + i] = (float)((double)a / d.A);
                planes = default_planes.data();
            }
            c.act_table.assign((size_t)d.A * hw * C, 0.f);
            for (int a = 0; a < d.A; ++a)
                for (int y = 0; y < d.h; ++y)
                    for (int x = 0; x < d.w; ++x)
                        for (int o = 0; o < C; ++o) {
                            double acc = 0;
                            for (int ch = 0; ch < d.ach; ++ch)
                                for (int ky = 0; ky < 3; ++ky)
                                    for (int kx = 0; kx < 3; ++kx) {
                                        int yy = y + ky - 1, xx = x + kx - 1;
                                        if (yy < 0 || yy >= d.h || xx < 0 || xx >= d.w) continue;
                                        acc += (double)w[((size_t)o * (C + d.ach) + C + ch) * 9 + ky * 3 + kx] *
                                               (double)planes[((size_t)a * d.ach + ch) * hw + yy * d.w + xx];
                                    }
                            double s = (double)gamma[o] / std::sqrt((double)var[o] + 1e-4);
                            c.act_table[((size_t)a * hw + y * d.w + x) * C + o] = (float)(acc * s);
                        }
        }
        add_conv(&dyn, std::move(c));
    }
    for (int i = 0; i < cfg.dynamics_blocks; ++i)
        add_block(g, &dyn, "dynamics_network.dynamics_resnet_body." + std::to_string(i), i, C, d.h, d.w);
    finish_tower(&dyn, C, d.h, d.w);

    // ---- prediction (models.cpp:122-166) ----
    Tower& pred = plan->prediction;
    pred = Tower();
    for (int i = 0; i < cfg.prediction_blocks; ++i)
        add_block(g, &pred, "prediction_network.prediction_resnet_body." + std::to_string(i), i, C, d.h, d.w);
    finish_tower(&pred, C, d.h, d.w);

    build_head(g, "dynamics_network.dynamics_reward_conv", "dynamics_network.dynamics_reward_mlp.dynamics_reward_mlp", C,
               cfg.reward_reduced_channels, d.hw(), cfg.reward_head_layers, cfg.n_reward_head_layers, d.Sr, &plan->reward_head);
    build_head(g, "prediction_network.prediction_policy_conv", "prediction_network.prediction_policy_mlp.prediction_policy_mlp", C,
               cfg.policy_reduced_channels, d.hw(), cfg.policy_head_layers, cfg.n_policy_head_layers, d.A, &plan->policy_head);
    build_head(g, "prediction_network.prediction_value_conv", "prediction_network.prediction_value_mlp.prediction_value_mlp", C,
               cfg.value_reduced_channels, d.hw(), cfg.value_head_layers, cfg.n_value_head_layers, d.Sv, &plan->value_head);
    if (!g.ok) { set_error("%s", g.missing.c_str()); return MZB_ERR_INVALID; }

    double fpred = tower_flops(pred) + head_flops(plan->policy_head) + head_flops(plan->value_head);
    plan->flops_initial = tower_flops(rep) + fpred;
    plan->flops_recurrent = tower_flops(dyn) + 2.0 * d.ach * C * 9.0 * d.hw() + head_flops(plan->reward_head) + fpred;
    return MZB_OK;
}

}  // namespace mzb
