// ref_driver — TEST INFRASTRUCTURE.  A thin command-line harness around the UNMODIFIED reference
// (compiled from /root/reference by oracle/ref_build/Makefile into oracle/_ref/).  It never
// links into the product.  Modes:
//   mcts   run algorithm::MCTS::run (muzero-cpp/mcts.cpp:34-87) against the synthetic evaluator "SE"
//          (defined in oracle/se_evaluator.h, shared with the C oracle) and dump results + the
//          (parent node, action) expansion trace in text with hex floats.
//   net    build a MuzeroNetwork (models.cpp:171-209), optionally randomise BN statistics, save it, load it
//          into a VPRNetModel (vprnet.cpp:30-50,71-74) and dump weights / inputs / outputs of
//          InitialInference + RecurrentInference (vprnet.cpp:77-189) as an MZBT tensor file.
//   replay run MCTS::run against an evaluator that replays recorded per-simulation network outputs
//          (MZBT file written by the engine) and dump results + trace (end-to-end decision parity).
//   bench  time MCTS::run over VPRNetEvaluator/DeviceManager on libtorch CPU (the reported CPU baseline).
//   encoder dump ValueEncoder support size / decode for given bounds (util.cpp:45-102).
//   stack  GameHistory::get_stacked_observations (types.cpp:16-42) on a synthetic game history.
//   select the action-selection rule of self_play.cpp:21-39 over libstdc++'s discrete_distribution.
#include <torch/torch.h>

#include <algorithm>
#include <atomic>
#include <cmath>
#include <random>
#include <chrono>
#include <cinttypes>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <iostream>
#include <map>
#include <memory>
#include <string>
#include <thread>
#include <vector>

// learn mode reads the gradients Learn() leaves on the module's parameters: VPRNetModel keeps the module private, so this test
// driver (and nothing else) opens the class up.  The reference sources stay unmodified; GCC lays the class out the same way.
#define private public
#include "muzero-cpp/vprnet.h"
#undef private
#include "muzero-cpp/config.h"
#include "muzero-cpp/device_manager.h"
#include "muzero-cpp/mcts.h"
#include "muzero-cpp/models.h"
#include "muzero-cpp/types.h"
#include "muzero-cpp/util.h"
#include "muzero-cpp/vprnet.h"
#include "muzero-cpp/vprnet_evaluator.h"

#include "../se_evaluator.h"
#include "../mzbt_io.h"

using namespace muzero_cpp;
using namespace muzero_cpp::types;
using muzero_cpp::muzero_config::MuZeroConfig;

// ---------------------------------------------------------------- argument parsing (key=value)
struct Args {
    std::map<std::string, std::string> kv;
    Args(int argc, char** argv) {
        for (int i = 2; i < argc; ++i) {
            std::string s(argv[i]);
            auto p = s.find('=');
            if (p == std::string::npos) { kv[s] = "1"; } else { kv[s.substr(0, p)] = s.substr(p + 1); }
        }
    }
    bool has(const std::string& k) const { return kv.count(k) > 0; }
    std::string str(const std::string& k, const std::string& d) const { return has(k) ? kv.at(k) : d; }
    int i(const std::string& k, int d) const { return has(k) ? std::stoi(kv.at(k)) : d; }
    double d(const std::string& k, double dflt) const { return has(k) ? std::stod(kv.at(k)) : dflt; }
};

static int g_plane_hw = 0;
static int g_num_actions = 1;
static int g_plane_hw_initial = 0;
// the constant-plane action encoding every shipped example uses
// (examples/connect4/muzero_connect4.cpp:103-108, examples/ale/muzero_ale.cpp:303-315)
static Observation encode_action_recurrent(Action a) { return Observation(g_plane_hw, (double)a / g_num_actions); }
static Observation encode_action_initial(Action a) { return Observation(g_plane_hw_initial, (double)a / g_num_actions); }

// Game presets = the shapes of BASELINE.json's configs (SURVEY.md §8 header).
static MuZeroConfig make_config(const Args& a) {
    MuZeroConfig c;
    std::string game = a.str("game", "c4d");
    c.seed = 0;
    c.path = a.str("path", "/tmp/mzb_ref_scratch");
    c.action_channels = 1;
    c.network_config.normalize_hidden_states = true;
    auto set_board = [&](int ch, int h, int w, int na, int players) {
        c.observation_shape = {ch, h, w};
        c.action_space.clear();
        for (int i = 0; i < na; ++i) c.action_space.push_back(i);
        c.num_players = players;
    };
    if (game == "c4d" || game == "c4r") {
        set_board(3, 6, 7, 7, 2);
        c.value_upperbound = 10; c.value_lowerbound = -10;
        c.min_reward = -10; c.max_reward = 10; c.min_value = -10; c.max_value = 10;
        if (game == "c4r") {
            c.network_config.resnet_channels = 64;
            c.network_config.representation_blocks = 3;
            c.network_config.dynamics_blocks = 3;
            c.network_config.prediction_blocks = 3;
        }
    } else if (game == "ttt") {
        set_board(3, 3, 3, 9, 2);
        c.value_upperbound = 10; c.value_lowerbound = -10;
        c.min_reward = -10; c.max_reward = 10; c.min_value = -10; c.max_value = 10;
    } else if (game == "ale") {
        set_board(1, 96, 96, 18, 1);
        c.stacked_observations = 6;
        c.network_config.downsample = true;
        c.network_config.reward_reduced_channels = 64;
        c.network_config.policy_reduced_channels = 64;
        c.network_config.value_reduced_channels = 64;
        c.network_config.reward_head_layers = {64, 64};
        c.network_config.policy_head_layers = {64, 64};
        c.network_config.value_head_layers = {64, 64};
        c.min_value = -21; c.max_value = 21; c.min_reward = -1; c.max_reward = 1;
    } else if (game == "tiny") {   // small golden-fixture config
        set_board(2, 3, 4, 5, 2);
        c.stacked_observations = 1;
        c.value_upperbound = 5; c.value_lowerbound = -5;
        c.min_reward = -3; c.max_reward = 3; c.min_value = -5; c.max_value = 5;
        c.network_config.resnet_channels = 16;
        c.network_config.representation_blocks = 1;
        c.network_config.dynamics_blocks = 1;
        c.network_config.prediction_blocks = 2;
        c.network_config.reward_reduced_channels = 4;
        c.network_config.policy_reduced_channels = 2;
        c.network_config.value_reduced_channels = 3;
        c.network_config.reward_head_layers = {8};
        c.network_config.policy_head_layers = {8, 6};
        c.network_config.value_head_layers = {};
    } else if (game == "tinyds") {   // small downsample fixture
        set_board(1, 32, 24, 6, 1);
        c.stacked_observations = 2;
        c.network_config.downsample = true;
        c.min_value = -21; c.max_value = 21; c.min_reward = -1; c.max_reward = 1;
        c.network_config.resnet_channels = 16;
        c.network_config.representation_blocks = 1;
        c.network_config.dynamics_blocks = 1;
        c.network_config.prediction_blocks = 1;
        c.network_config.reward_reduced_channels = 4;
        c.network_config.policy_reduced_channels = 4;
        c.network_config.value_reduced_channels = 4;
        c.network_config.reward_head_layers = {8};
        c.network_config.policy_head_layers = {8};
        c.network_config.value_head_layers = {8};
    } else {
        std::cerr << "unknown game " << game << std::endl;
        std::exit(2);
    }
    // overrides
    if (a.has("channels")) c.network_config.resnet_channels = a.i("channels", 128);
    if (a.has("blocks")) {
        c.network_config.representation_blocks = c.network_config.dynamics_blocks =
            c.network_config.prediction_blocks = a.i("blocks", 4);
    }
    if (a.has("no_normalize")) c.network_config.normalize_hidden_states = false;
    if (a.has("no_contractive")) c.use_contractive_mapping = false;
    c.num_simulations = a.i("sims", 50);
    c.discount = a.d("discount", 0.997);
    c.pb_c_base = a.d("pb_c_base", 19652);
    c.pb_c_init = a.d("pb_c_init", 1.25);
    c.dirichlet_alpha = a.d("alpha", 0.3);
    c.dirichlet_epsilon = a.d("eps", 0.25);
    if (a.has("players")) c.num_players = a.i("players", 2);
    if (a.has("unbounded")) { c.value_upperbound = NINF_D; c.value_lowerbound = INF_D; }
    if (a.has("lower")) c.value_lowerbound = a.d("lower", 0);
    if (a.has("upper")) c.value_upperbound = a.d("upper", 1);
    if (a.has("actions")) {
        c.action_space.clear();
        for (int i = 0; i < a.i("actions", 7); ++i) c.action_space.push_back(i);
    }
    ObservationShape enc = model::RepresentationNetworkImpl::encoded_state_shape(
        {c.network_config.resnet_channels, c.observation_shape.h, c.observation_shape.w},
        c.network_config.downsample);
    g_plane_hw = enc.h * enc.w;
    g_plane_hw_initial = c.observation_shape.h * c.observation_shape.w;
    g_num_actions = (int)c.action_space.size();
    c.action_representation_initial = encode_action_initial;
    c.action_representation_recurrent = encode_action_recurrent;
    c.visit_softmax_temperature = [](int) { return 1.0; };
    return c;
}

// ---------------------------------------------------------------- synthetic evaluator (mode mcts)
// Hidden state = {node index (expansion order), content id}; see oracle/se_evaluator.h.
struct TraceRec { int parent; int action; };
class SEEvaluator : public Evaluator {
public:
    SEEvaluator(int num_actions, double vlo, double vhi, double rlo, double rhi)
        : na_(num_actions), vlo_(vlo), vhi_(vhi), rlo_(rlo), rhi_(rhi) {}
    VPRNetModel::InferenceOutputs InitialInference(const Observation& obs) override {
        uint32_t cid = se_initial_cid((uint32_t)obs[0]);
        calls_ = 0;
        trace.clear();
        return make(cid, 0);
    }
    VPRNetModel::InferenceOutputs RecurrentInference(const Observation& hidden, Action action) override {
        uint32_t cid = se_child_cid((uint32_t)hidden[1], (uint32_t)action);
        ++calls_;
        trace.push_back({(int)hidden[0], action});
        return make(cid, calls_);
    }
    std::vector<TraceRec> trace;

private:
    VPRNetModel::InferenceOutputs make(uint32_t cid, int node) {
        VPRNetModel::InferenceOutputs o;
        o.value = se_value(cid, vlo_, vhi_);
        o.reward = se_reward(cid, rlo_, rhi_);
        o.policy.resize(na_);
        se_policy(cid, na_, o.policy.data());
        o.encoded_state = {(float)node, (float)cid};
        return o;
    }
    int na_;
    double vlo_, vhi_, rlo_, rhi_;
    int calls_ = 0;
};

static void print_hexd(FILE* f, const std::vector<double>& v) {
    for (double x : v) fprintf(f, " %a", x);
}

static void dump_mcts_result(FILE* f, int tree, int run, const MCTSReturn& r, const std::vector<TraceRec>& trace) {
    fprintf(f, "tree %d run %d root_value %a depth %d nchild %d\n", tree, run, r.root_value, r.max_tree_depth,
            (int)r.child_actions.size());
    fprintf(f, "actions");
    for (int a : r.child_actions) fprintf(f, " %d", a);
    fprintf(f, "\nrelvisits");
    print_hexd(f, r.children_relative_visit);
    fprintf(f, "\npolicy");
    print_hexd(f, r.emperical_policy);
    fprintf(f, "\ntrace");
    for (auto& t : trace) fprintf(f, " %d:%d", t.parent, t.action);
    fprintf(f, "\n");
}

// legal-action pattern shared with the oracle and the engine tests: see se_legal_mask().
static std::vector<Action> legal_for(uint32_t key, int na, int mode) {
    std::vector<Action> legal;
    uint32_t mask = se_legal_mask(key, na, mode);
    for (int a = 0; a < na; ++a) if (mask & (1u << a)) legal.push_back(a);
    return legal;
}

static int mode_mcts(const Args& a) {
    MuZeroConfig c = make_config(a);
    int T = a.i("trees", 4), runs = a.i("runs", 1), seed = a.i("seed", 0), legal_mode = a.i("legal_mode", 0);
    bool noise = a.i("noise", 1) != 0;
    double vlo = a.d("se_vlo", -1), vhi = a.d("se_vhi", 1), rlo = a.d("se_rlo", -0.5), rhi = a.d("se_rhi", 0.5);
    FILE* f = a.has("out") ? fopen(a.str("out", "").c_str(), "w") : stdout;
    int na = (int)c.action_space.size();
    for (int t = 0; t < T; ++t) {
        auto ev = std::make_shared<SEEvaluator>(na, vlo, vhi, rlo, rhi);
        // one MCTS object per tree slot, seeded seed+tree (self_play.cpp:232 uses config.seed + actor_num)
        algorithm::MCTS mcts(c, seed + t, ev);
        for (int r = 0; r < runs; ++r) {
            uint32_t key = (uint32_t)(t * 7919 + r * 104729 + 1) & 0xFFFFFF;
            Observation obs = {(float)key};
            std::vector<Action> legal = legal_for(key, na, legal_mode);
            Player to_play = (t + r) % c.num_players;
            MCTSReturn ret = mcts.run(obs, legal, to_play, noise);
            dump_mcts_result(f, t, r, ret, ev->trace);
        }
    }
    if (f != stdout) fclose(f);
    return 0;
}

// ---------------------------------------------------------------- mode net
static void add_f64(mzbt::Writer& w, const std::string& n, const std::vector<int64_t>& dims, const std::vector<double>& v) {
    w.add(n, 1, dims, v.data());
}
static void add_tensor_f32(mzbt::Writer& w, const std::string& n, torch::Tensor t) {
    t = t.detach().to(torch::kCPU).to(torch::kFloat).contiguous();
    std::vector<int64_t> dims(t.sizes().begin(), t.sizes().end());
    w.add(n, 0, dims, t.data_ptr<float>());
}
static void dump_outputs(mzbt::Writer& w, const std::string& pre, const std::vector<VPRNetModel::InferenceOutputs>& o) {
    int64_t B = (int64_t)o.size(), A = (int64_t)o[0].policy.size(), S = (int64_t)o[0].encoded_state.size();
    std::vector<double> v(B), r(B), p(B * A);
    std::vector<float> h(B * S);
    for (int64_t b = 0; b < B; ++b) {
        v[b] = o[b].value; r[b] = o[b].reward;
        for (int64_t i = 0; i < A; ++i) p[b * A + i] = o[b].policy[i];
        std::memcpy(&h[b * S], o[b].encoded_state.data(), S * sizeof(float));
    }
    add_f64(w, pre + ".value", {B}, v);
    add_f64(w, pre + ".reward", {B}, r);
    add_f64(w, pre + ".policy", {B, A}, p);
    w.add(pre + ".hidden", 0, {B, S}, h.data());
}

static void randomize_bn(model::MuzeroNetwork& net) {
    torch::NoGradGuard ng;
    for (auto& kv : net->named_buffers()) {
        const std::string& k = kv.key();
        if (k.size() > 12 && k.compare(k.size() - 12, 12, "running_mean") == 0) kv.value().copy_(torch::randn_like(kv.value()) * 0.3);
        if (k.size() > 11 && k.compare(k.size() - 11, 11, "running_var") == 0) kv.value().copy_(torch::rand_like(kv.value()) + 0.5);
    }
    for (auto& kv : net->named_parameters()) {
        const std::string& k = kv.key();
        bool is_bn = k.find("_bn") != std::string::npos;
        if (!is_bn) continue;
        if (k.size() > 6 && k.compare(k.size() - 6, 6, "weight") == 0) kv.value().copy_(torch::rand_like(kv.value()) + 0.5);
        if (k.size() > 4 && k.compare(k.size() - 4, 4, "bias") == 0) kv.value().copy_(torch::randn_like(kv.value()) * 0.2);
    }
}

static void dump_weights(mzbt::Writer& w, model::MuzeroNetwork& net) {
    for (auto& kv : net->named_parameters()) add_tensor_f32(w, "w." + kv.key(), kv.value());
    for (auto& kv : net->named_buffers()) {
        if (kv.key().find("num_batches_tracked") != std::string::npos) continue;
        add_tensor_f32(w, "w." + kv.key(), kv.value());
    }
}

static int mode_net(const Args& a) {
    MuZeroConfig c = make_config(a);
    int B = a.i("batch", 4), seed = a.i("seed", 0);
    torch::manual_seed(seed);
    model::MuzeroNetwork net(c);
    if (a.has("ckpt")) { torch::load(net, a.str("ckpt", ""), torch::Device(torch::kCPU)); }   // the shipped archive was saved from cuda:0
    else if (a.i("randomize_bn", 1)) { randomize_bn(net); }
    net->eval();

    // Route the same weights through VPRNetModel (vprnet.cpp:58-74 needs both files to exist).
    VPRNetModel vm(c, "cpu");
    std::string base = vm.SaveCheckpoint(-7);
    torch::save(net, base + ".pt");
    vm.LoadCheckpoint(-7);
    if (a.has("ckpt_out")) {   // keep the archive torch::save wrote (the file format VPRNetModel::LoadCheckpoint reads) as a fixture
        std::ifstream src(base + ".pt", std::ios::binary);
        std::ofstream dst(a.str("ckpt_out", ""), std::ios::binary);
        dst << src.rdbuf();
    }

    mzbt::Writer w;
    if (!a.has("no_weights")) dump_weights(w, net);   // no_weights: the fixture refers to the checkpoint file itself
    int A = (int)c.action_space.size();
    int cin = net->get_initial_inference_channels();
    int flat = cin * c.observation_shape.h * c.observation_shape.w;
    torch::Tensor obs = torch::rand({B, flat});
    if (a.i("binary_obs", 0)) obs = (obs < 0.3).to(torch::kFloat);
    if (a.has("obs_in")) {   // observations given by the caller (real game positions): MZBT file with a float tensor "in.obs" [B, flat]
        std::vector<mzbt::Tensor> ts;
        if (!mzbt::load(a.str("obs_in", ""), ts)) { std::cerr << "cannot read obs_in" << std::endl; return 1; }
        const mzbt::Tensor* given = nullptr;
        for (auto& t : ts) if (t.name == "in.obs") given = &t;
        if (!given || given->dims.size() != 2 || given->dims[1] != flat) { std::cerr << "obs_in: no in.obs of width " << flat << std::endl; return 1; }
        B = (int)given->dims[0];
        obs = torch::from_blob((void*)given->as<float>(), {B, flat}, torch::kFloat).clone();
    }
    std::vector<VPRNetModel::InitialInferenceInputs> in0(B);
    for (int b = 0; b < B; ++b) in0[b].stacked_observation = Observation(obs[b].data_ptr<float>(), obs[b].data_ptr<float>() + flat);
    auto o0 = vm.InitialInference(in0);
    std::vector<int32_t> a1(B), a2(B);
    for (int b = 0; b < B; ++b) { a1[b] = (int)(se_mix(seed * 977 + b) % A); a2[b] = (int)(se_mix(seed * 3571 + b + 17) % A); }
    std::vector<VPRNetModel::RecurrentInferenceInputs> in1(B), in2(B);
    for (int b = 0; b < B; ++b) { in1[b].action = a1[b]; in1[b].encoded_state = o0[b].encoded_state; }
    auto o1 = vm.RecurrentInference(in1);
    for (int b = 0; b < B; ++b) { in2[b].action = a2[b]; in2[b].encoded_state = o1[b].encoded_state; }
    auto o2 = vm.RecurrentInference(in2);

    add_tensor_f32(w, "in.obs", obs);
    w.add("in.a1", 2, {B}, a1.data());
    w.add("in.a2", 2, {B}, a2.data());
    dump_outputs(w, "init", o0);
    dump_outputs(w, "rec1", o1);
    dump_outputs(w, "rec2", o2);
    {   // raw logits straight from the module, for debugging the heads
        torch::NoGradGuard ng;
        auto enc = net->get_encoded_observation_shape();
        auto r0 = net->initial_inference(obs.reshape({B, cin, c.observation_shape.h, c.observation_shape.w}));
        add_tensor_f32(w, "dbg.init.value_logits", r0.value);
        add_tensor_f32(w, "dbg.init.policy_logits", r0.policy_logits);
        torch::Tensor act = torch::empty({B, 1, enc.h, enc.w});
        for (int b = 0; b < B; ++b) act[b].fill_((float)((double)a1[b] / A));
        auto r1 = net->recurrent_inference(r0.encoded_state, act);
        add_tensor_f32(w, "dbg.rec1.value_logits", r1.value);
        add_tensor_f32(w, "dbg.rec1.reward_logits", r1.reward);
        add_tensor_f32(w, "dbg.rec1.policy_logits", r1.policy_logits);
    }
    std::vector<int32_t> meta = {cin, c.observation_shape.h, c.observation_shape.w, A,
                                 net->get_encoded_observation_shape().c, net->get_encoded_observation_shape().h,
                                 net->get_encoded_observation_shape().w,
                                 util::ValueEncoder::get_support_size(c.min_value, c.max_value, c.use_contractive_mapping),
                                 util::ValueEncoder::get_support_size(c.min_reward, c.max_reward, c.use_contractive_mapping)};
    w.add("meta", 2, {(int64_t)meta.size()}, meta.data());
    if (!w.save(a.str("out", "/tmp/mzb_ref_net.mzbt"))) { std::cerr << "cannot write output" << std::endl; return 1; }
    return 0;
}

// ---------------------------------------------------------------- mode learn
// One call of the UNMODIFIED VPRNetModel::Learn (vprnet.cpp:192-311) on a seeded types::Batch; dumps the batch, the LossInfo it
// returns (losses are computed before the optimiser step) and the BatchNorm running statistics after the call (they are moved by
// the forward passes alone: momentum 0.01, model_layers.cpp:36-42).  Weights: seeded random init (dumped) or ckpt=<module archive>.
static int mode_learn(const Args& a) {
    MuZeroConfig c = make_config(a);
    const int B = a.i("batch", 6), seed = a.i("seed", 0);
    c.num_unroll_steps = a.i("unroll", 5);
    c.value_loss_weight = a.d("value_loss_weight", 0.25);
    c.network_config.learning_rate = a.d("lr", c.network_config.learning_rate);
    c.network_config.l2_weight_decay = a.d("l2", c.network_config.l2_weight_decay);
    torch::manual_seed(seed);
    model::MuzeroNetwork net(c);
    if (a.has("ckpt")) { torch::load(net, a.str("ckpt", ""), torch::Device(torch::kCPU)); }
    else { randomize_bn(net); }
    VPRNetModel vm(c, "cpu");
    std::string base = vm.SaveCheckpoint(-7);
    torch::save(net, base + ".pt");
    vm.LoadCheckpoint(-7);
    mzbt::Writer w;
    if (!a.has("no_weights")) dump_weights(w, net);
    const int K1 = c.num_unroll_steps + 1, A = (int)c.action_space.size();
    const int cin = net->get_initial_inference_channels();
    const int flat = cin * c.observation_shape.h * c.observation_shape.w;
    Batch batch;
    batch.num_samples = B;
    std::mt19937 rng(seed * 7919 + 3);
    std::uniform_real_distribution<double> u(0.0, 1.0);
    for (int b = 0; b < B; ++b) {
        batch.indices.push_back(b);
        batch.priorities.push_back(0.2 + 0.8 * u(rng));
        batch.gradient_scale.push_back(1.0 / c.num_unroll_steps);
        for (int i = 0; i < flat; ++i) batch.stacked_observations.push_back(a.i("binary_obs", 0) ? (u(rng) < 0.3 ? 1.f : 0.f) : (float)u(rng));
        for (int k = 0; k < K1; ++k) {
            batch.actions.push_back((int)(rng() % A));
            batch.target_values.push_back(c.min_value + (c.max_value - c.min_value) * u(rng));
            batch.target_rewards.push_back(c.min_reward + (c.max_reward - c.min_reward) * u(rng));
            double sum = 0;
            std::vector<double> p(A);
            for (int i = 0; i < A; ++i) { p[i] = u(rng) < 0.3 ? 0.0 : u(rng); sum += p[i]; }
            for (int i = 0; i < A; ++i) batch.target_policies.push_back(sum > 0 ? p[i] / sum : 1.0 / A);
        }
    }
    std::vector<int32_t> acts(batch.actions.begin(), batch.actions.end());
    w.add("in.obs", 0, {B, flat}, batch.stacked_observations.data());
    w.add("in.actions", 2, {B, K1}, acts.data());
    add_f64(w, "in.target_values", {B, K1}, batch.target_values);
    add_f64(w, "in.target_rewards", {B, K1}, batch.target_rewards);
    add_f64(w, "in.target_policies", {B, K1, A}, batch.target_policies);
    add_f64(w, "in.gradient_scale", {B}, batch.gradient_scale);
    add_f64(w, "in.priorities", {B}, batch.priorities);
    // steps=N: N Learn calls on the same batch (the optimizer's moments and the BatchNorm buffers move from call to call)
    const int n_steps = a.i("steps", 1);
    VPRNetModel::LossInfo li = vm.Learn(batch);
    std::vector<double> all_losses = {li.total_loss, li.value, li.policy, li.reward};
    for (int it = 1; it < n_steps; ++it) {
        li = vm.Learn(batch);
        for (double v : {li.total_loss, li.value, li.policy, li.reward}) all_losses.push_back(v);
    }
    add_f64(w, "out.losses", {4}, {li.total_loss, li.value, li.policy, li.reward});
    add_f64(w, "out.all_losses", {n_steps, 4}, all_losses);
    add_f64(w, "out.errors", {(int64_t)li.errors.size()}, li.errors);
    if (a.has("grads")) {
        // what backward() left on the parameters in the LAST call, and the parameters after the last optimizer step.
        // grads=full: every tensor; grads=digest: per tensor {sum g^2, sum g_i cos(0.731 i)} and {the same of the parameter}
        const bool full = a.str("grads", "full") == "full";
        for (auto& kv : vm.model_->named_parameters()) {
            torch::Tensor g = kv.value().grad();
            if (!g.defined()) continue;
            torch::Tensor p = kv.value().detach();
            if (full) {
                add_tensor_f32(w, "grad." + kv.key(), g);
                add_tensor_f32(w, "param." + kv.key(), p);
            } else {
                torch::Tensor probe = torch::cos(torch::arange(g.numel(), torch::kFloat64) * 0.731);
                auto dig = [&](const torch::Tensor& t) {
                    torch::Tensor f = t.contiguous().view({-1}).to(torch::kFloat64);
                    return std::vector<double>{(f * f).sum().item<double>(), (f * probe).sum().item<double>()};
                };
                add_f64(w, "graddigest." + kv.key(), {2}, dig(g));
                add_f64(w, "paramdigest." + kv.key(), {2}, dig(p));
            }
        }
        std::vector<double> hp = {c.network_config.learning_rate, c.network_config.l2_weight_decay};
        add_f64(w, "meta.optimizer", {2}, hp);
    }
    // BatchNorm buffers after the call: through the model's own checkpoint
    std::string after = vm.SaveCheckpoint(-8);
    model::MuzeroNetwork net2(c);
    torch::load(net2, after + ".pt");
    for (auto& kv : net2->named_buffers()) {
        const std::string& k = kv.key();
        if (k.find("running_") != std::string::npos && (!a.has("post_filter") || k.find(a.str("post_filter", "")) != std::string::npos))
            add_tensor_f32(w, "post." + k, kv.value());
    }
    std::vector<int32_t> meta = {cin, c.observation_shape.h, c.observation_shape.w, A, c.num_unroll_steps};
    w.add("meta", 2, {(int64_t)meta.size()}, meta.data());
    if (!w.save(a.str("out", "/tmp/mzb_ref_learn.mzbt"))) { std::cerr << "cannot write output" << std::endl; return 1; }
    printf("learn: total %.9g value %.9g policy %.9g reward %.9g\n", li.total_loss, li.value, li.policy, li.reward);
    return 0;
}

// ---------------------------------------------------------------- mode replay
// Replays network outputs recorded by the engine, so the reference's search runs on "identical network
// outputs" (north star) and its decisions can be compared with the engine's, simulation by simulation.
class ReplayEvaluator : public Evaluator {
public:
    ReplayEvaluator(int na, int sims) : na_(na), sims_(sims) {}
    void set(const double* v, const double* r, const double* p) { v_ = v; r_ = r; p_ = p; }
    VPRNetModel::InferenceOutputs InitialInference(const Observation&) override { calls_ = 0; trace.clear(); return make(0); }
    VPRNetModel::InferenceOutputs RecurrentInference(const Observation& hidden, Action action) override {
        ++calls_;
        trace.push_back({(int)hidden[0], action});
        return make(calls_);
    }
    std::vector<TraceRec> trace;
private:
    VPRNetModel::InferenceOutputs make(int k) {
        VPRNetModel::InferenceOutputs o;
        o.value = v_[k]; o.reward = r_[k];
        o.policy.assign(p_ + (size_t)k * na_, p_ + (size_t)(k + 1) * na_);
        o.encoded_state = {(float)k};
        return o;
    }
    int na_, sims_, calls_ = 0;
    const double *v_ = nullptr, *r_ = nullptr, *p_ = nullptr;
};

static const mzbt::Tensor* find_t(const std::vector<mzbt::Tensor>& ts, const std::string& n) {
    for (auto& t : ts) if (t.name == n) return &t;
    std::cerr << "missing tensor " << n << std::endl;
    std::exit(3);
}

static int mode_replay(const Args& a) {
    MuZeroConfig c = make_config(a);
    std::vector<mzbt::Tensor> ts;
    if (!mzbt::load(a.str("in", ""), ts)) { std::cerr << "cannot read replay file" << std::endl; return 1; }
    // value/reward [T, sims+1] f64, policy [T, sims+1, A] f64, legal_mask [T] i32, to_play [T] i32, noise [T] i32, seeds [T] i32
    const mzbt::Tensor* tv = find_t(ts, "value"); const mzbt::Tensor* tr = find_t(ts, "reward");
    const mzbt::Tensor* tp = find_t(ts, "policy"); const mzbt::Tensor* tl = find_t(ts, "legal_mask");
    const mzbt::Tensor* tt = find_t(ts, "to_play"); const mzbt::Tensor* tn = find_t(ts, "noise");
    const mzbt::Tensor* tsd = find_t(ts, "seeds");
    int T = (int)tv->dims[0], n1 = (int)tv->dims[1], na = (int)c.action_space.size();
    c.num_simulations = n1 - 1;
    FILE* f = a.has("out") ? fopen(a.str("out", "").c_str(), "w") : stdout;
    for (int t = 0; t < T; ++t) {
        auto ev = std::make_shared<ReplayEvaluator>(na, n1 - 1);
        ev->set(tv->as<double>() + (size_t)t * n1, tr->as<double>() + (size_t)t * n1, tp->as<double>() + (size_t)t * n1 * na);
        algorithm::MCTS mcts(c, tsd->as<int32_t>()[t], ev);
        std::vector<Action> legal;
        uint32_t mask = (uint32_t)tl->as<int32_t>()[t];
        for (int i = 0; i < na; ++i) if (mask & (1u << i)) legal.push_back(i);
        MCTSReturn ret = mcts.run({0.f}, legal, tt->as<int32_t>()[t], tn->as<int32_t>()[t] != 0);
        dump_mcts_result(f, t, 0, ret, ev->trace);
    }
    if (f != stdout) fclose(f);
    return 0;
}

// ---------------------------------------------------------------- mode bench (reported CPU baseline)
// The reference's own path: actor threads each own an MCTS (self_play.cpp:228-238) sharing one
// VPRNetEvaluator (vprnet_evaluator.cpp:25-178) over a DeviceManager of libtorch-CPU VPRNetModels
// (muzero.cpp:60-63,161-163) — i.e. --devices=cpu,cpu semantics.
static std::vector<int> int_list(const std::string& s) {
    std::vector<int> v;
    size_t at = 0;
    while (at < s.size()) {
        size_t c = s.find(',', at);
        if (c == std::string::npos) c = s.size();
        if (c > at) v.push_back(std::stoi(s.substr(at, c - at)));
        at = c + 1;
    }
    return v;
}

// One (actors, evaluator threads) arrangement: fresh DeviceManager + VPRNetEvaluator, `warmup` untimed and `steps` timed
// steps of `moves` searches per actor thread.  Returns the seconds of every timed step.
static std::vector<double> bench_combo(const MuZeroConfig& c, int actors, int eval_threads, int devices, int seed, int flat,
                                       int moves, int warmup, int steps) {
    DeviceManager dm;
    for (int d = 0; d < devices; ++d) dm.AddDevice(VPRNetModel(c, "cpu"));
    auto ev = std::make_shared<VPRNetEvaluator>(&dm, actors, eval_threads, actors, eval_threads);   // evaluator batch = actors (README.md:87)
    std::vector<Action> legal = c.action_space;
    auto body = [&](int id, int nmoves) {
        algorithm::MCTS mcts(c, seed + id, ev);
        std::mt19937 rng(seed + id);
        std::uniform_real_distribution<float> u(0.f, 1.f);
        Observation obs(flat);
        for (int m = 0; m < nmoves; ++m) {
            for (auto& x : obs) x = u(rng);
            MCTSReturn r = mcts.run(obs, legal, m % c.num_players, true);
            (void)r;
        }
    };
    std::vector<double> secs;
    for (int s = 0; s < warmup + steps; ++s) {
        auto t0 = std::chrono::steady_clock::now();
        std::vector<std::thread> th;
        for (int i = 0; i < actors; ++i) th.emplace_back(body, i, s < warmup ? 1 : moves);
        for (auto& t : th) t.join();
        if (s >= warmup) secs.push_back(std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count());
    }
    return secs;
}

static int mode_bench(const Args& a) {
    MuZeroConfig c = make_config(a);
    int moves = a.i("moves", 2), devices = a.i("devices", 2), seed = a.i("seed", 0);
    int steps = a.i("steps", 1), warmup = a.i("warmup", 1);
    std::vector<int> actor_list = int_list(a.str("actors", "4")), et_list = int_list(a.str("eval_threads", "1"));
    if (a.has("torch_threads")) torch::set_num_threads(a.i("torch_threads", 1));
    torch::manual_seed(seed);
    model::MuzeroNetwork probe(c);
    int flat = probe->get_initial_inference_channels() * c.observation_shape.h * c.observation_shape.w;
    // sweep (actors in {4, cores} with evaluator batch = actors; here also 1 or 2 evaluator threads): one short
    // timed step per arrangement, the best one then runs the requested steps
    int best_actors = actor_list[0], best_et = et_list[0];
    double best_rate = -1;
    std::string sweep = "[";
    if (actor_list.size() * et_list.size() > 1) {
        for (int ac : actor_list)
            for (int et : et_list) {
                std::vector<double> s = bench_combo(c, ac, et, devices, seed, flat, 1, 1, 1);
                double rate = (double)ac * c.num_simulations / s[0];
                char buf[160];
                snprintf(buf, sizeof buf, "%s{\"actors\": %d, \"eval_threads\": %d, \"sims_per_s\": %.3f}", sweep.size() > 1 ? ", " : "", ac, et, rate);
                sweep += buf;
                if (rate > best_rate) { best_rate = rate; best_actors = ac; best_et = et; }
            }
    }
    sweep += "]";
    std::vector<double> secs = bench_combo(c, best_actors, best_et, devices, seed, flat, moves, warmup, steps);
    double total = 0;
    for (double x : secs) total += x;
    long sims = (long)best_actors * moves * c.num_simulations * steps;
    printf("{\"mode\": \"bench\", \"game\": \"%s\", \"sims\": %ld, \"seconds\": %.6f, \"sims_per_s\": %.3f, \"actors\": %d, "
           "\"batch\": %d, \"eval_threads\": %d, \"devices\": %d, \"moves\": %d, \"steps\": %d, \"warmup\": %d, \"torch_threads\": %d, \"hw_threads\": %u, "
           "\"sweep\": %s, \"step_seconds\": [",
           a.str("game", "c4d").c_str(), sims, total, sims / total, best_actors, best_actors, best_et, devices, moves, steps, warmup,
           torch::get_num_threads(), std::thread::hardware_concurrency(), sweep.c_str());
    for (size_t i = 0; i < secs.size(); ++i) printf("%s%.6f", i ? ", " : "", secs[i]);
    printf("]}\n");
    return 0;
}

// VPRNetModel::RecurrentInference alone (states/s), vprnet.cpp:131-189
static int mode_bench_rec(const Args& a) {
    MuZeroConfig c = make_config(a);
    int B = a.i("batch", 64), iters = a.i("iters", 3);
    if (a.has("torch_threads")) torch::set_num_threads(a.i("torch_threads", 1));
    torch::manual_seed(0);
    VPRNetModel vm(c, "cpu");
    model::MuzeroNetwork probe(c);
    auto enc = probe->get_encoded_observation_shape();
    int S = enc.c * enc.h * enc.w, A = (int)c.action_space.size();
    std::vector<VPRNetModel::RecurrentInferenceInputs> in(B);
    std::mt19937 rng(0);
    std::uniform_real_distribution<float> u(0.f, 1.f);
    for (int b = 0; b < B; ++b) { in[b].action = b % A; in[b].encoded_state.resize(S); for (auto& x : in[b].encoded_state) x = u(rng); }
    vm.RecurrentInference(in);
    auto t0 = std::chrono::steady_clock::now();
    for (int i = 0; i < iters; ++i) vm.RecurrentInference(in);
    double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    printf("{\"mode\": \"bench_rec\", \"game\": \"%s\", \"batch\": %d, \"iters\": %d, \"seconds\": %.6f, \"states_per_s\": %.3f, "
           "\"torch_threads\": %d, \"hw_threads\": %u}\n",
           a.str("game", "c4d").c_str(), B, iters, sec, (double)B * iters / sec, torch::get_num_threads(),
           std::thread::hardware_concurrency());
    return 0;
}

// ---------------------------------------------------------------- mode encoder
static int mode_encoder(const Args& a) {
    double lo = a.d("min", -10), hi = a.d("max", 10);
    bool contractive = !a.has("no_contractive");
    util::ValueEncoder enc(lo, hi, contractive);
    int S = enc.get_support_size();
    printf("support %d\n", S);
    // decode of one-hot logits(0 / -inf approximated by log(eps)) and of a few seeded rows
    torch::manual_seed(a.i("seed", 0));
    int rows = a.i("rows", 8);
    torch::Tensor logits = torch::randn({rows, S}) * 2.0;
    torch::Tensor dec = enc.decode(logits).to(torch::kDouble);
    for (int r = 0; r < rows; ++r) {
        printf("row");
        for (int s = 0; s < S; ++s) printf(" %a", (double)logits[r][s].item<float>());
        printf(" -> %a\n", dec[r][0].item<double>());
    }
    return 0;
}

// stack: GameHistory::get_stacked_observations (types.cpp:16-42), the UNMODIFIED reference function, on a synthetic game:
// observation t = c*h*w values (t+1) + 0.001*k (k = flat index), action_history = {0, a_1, a_2, ...} with a_t = (7*t + 3) % A
// (self_play.cpp:91-92,158-159 build the two vectors exactly so), action planes = constant a / A (the shipped examples').
// Prints the stacked observation for every step of the game, then for the python-style negative index -1.
static int mode_stack(const Args& a) {
    const int c = a.i("c", 2), h = a.i("h", 3), w = a.i("w", 4), ach = a.i("ach", 1), S = a.i("stack", 3), steps = a.i("steps", 6), A = a.i("actions", 5);
    g_plane_hw_initial = ach * h * w;
    g_num_actions = A;
    types::GameHistory hist;
    types::ObservationShape shape{c, h, w};
    for (int t = 0; t < steps; ++t) {
        Observation o(c * h * w);
        for (int k = 0; k < c * h * w; ++k) o[k] = (float)(t + 1) + 0.001f * k;
        hist.action_history.push_back(t == 0 ? 0 : (7 * t + 3) % A);
        hist.observation_history.push_back(o);
    }
    printf("shape c=%d h=%d w=%d ach=%d stack=%d steps=%d actions=%d\n", c, h, w, ach, S, steps, A);
    for (int t = -1; t < steps; ++t) {
        Observation st = hist.get_stacked_observations(t, S, shape, ach, encode_action_initial);
        printf("step %d n=%zu", t, st.size());
        for (float v : st) printf(" %a", (double)v);
        printf("\n");
    }
    return 0;
}

// select: the action-selection rule of self_play.cpp:21-39 driven by libstdc++'s OWN std::discrete_distribution / std::mt19937
// (the third-party arithmetic the reference function delegates to; the function itself sits in an anonymous namespace and
// cannot be linked, so its few lines are restated here around the real library classes).  One rng per run, `draws` selections
// from it over random visit distributions: prints the inputs and the chosen action of every draw.
static int mode_select(const Args& a) {
    const int seed = a.i("seed", 0), draws = a.i("draws", 20), A = a.i("actions", 7);
    const double temperature = a.d("temperature", 1.0);
    std::mt19937 rng(seed);
    std::mt19937 gen(seed * 7919 + 13);   // input generator, separate from the rng under test
    printf("select seed=%d draws=%d actions=%d temperature=%a\n", seed, draws, A, temperature);
    for (int d = 0; d < draws; ++d) {
        const int n = 1 + (int)(gen() % A);
        std::vector<int> acts(A);
        for (int i = 0; i < A; ++i) acts[i] = i;
        for (int i = A - 1; i > 0; --i) std::swap(acts[i], acts[gen() % (i + 1)]);
        acts.resize(n);
        std::vector<int> visits(n);
        int total = 0;
        for (int i = 0; i < n; ++i) { visits[i] = (int)(gen() % 40); total += visits[i]; }
        if (total == 0) { visits[0] = 1; total = 1; }
        std::vector<double> rel(n);
        for (int i = 0; i < n; ++i) rel[i] = (double)visits[i] / total;
        int chosen;
        if (temperature <= 0) {
            chosen = acts[std::distance(rel.begin(), std::max_element(rel.begin(), rel.end()))];
        } else {
            std::vector<double> wts = rel;
            for (auto& v : wts) v = std::pow(v, 1.0 / temperature);
            std::discrete_distribution<> dist(wts.begin(), wts.end());
            chosen = acts[dist(rng)];
        }
        printf("draw n=%d", n);
        for (int i = 0; i < n; ++i) printf(" %d:%a", acts[i], rel[i]);
        printf(" -> %d\n", chosen);
    }
    return 0;
}

int main(int argc, char** argv) {
    if (argc < 2) { std::cerr << "usage: ref_driver <mcts|net|replay|bench|bench_rec|encoder> key=value ..." << std::endl; return 2; }
    Args a(argc, argv);
    std::string mode = argv[1];
    if (mode == "mcts") return mode_mcts(a);
    if (mode == "net") return mode_net(a);
    if (mode == "replay") return mode_replay(a);
    if (mode == "learn") return mode_learn(a);
    if (mode == "bench") return mode_bench(a);
    if (mode == "bench_rec") return mode_bench_rec(a);
    if (mode == "encoder") return mode_encoder(a);
    if (mode == "stack") return mode_stack(a);
    if (mode == "select") return mode_select(a);
    std::cerr << "unknown mode " << mode << std::endl;
    return 2;
}
