// muzero_cpp_mzb200.hpp — the reference's OWN names for the search hot path, backed by libmzb200.so.
//
// A muzero-cpp translation unit that today includes mcts.h / vprnet.h / vprnet_evaluator.h / device_manager.h / config.h /
// abstract_game.h for the self-play search (self_play.cpp, muzero.cpp:60-63,82,109-112,161-163, tests/mcts_test.cpp) compiles
// against this header unchanged: same namespaces, class names, constructor and method signatures, value-semantics and
// "no error codes" behaviour (a failing engine call prints mzb_last_error() and aborts).  Header-only, C++17, no torch, no abseil.
//
//   reference (file:line)                                   here
//   ------------------------------------------------------  ---------------------------------------------------------------
//   types::* (types.h:18-78)                                 muzero_cpp::types::* (aliases, ObservationShape, StepReturn, MCTSReturn)
//   muzero_config::MuZeroConfig (config.h:15-117)            same fields and defaults (the learner / replay fields are carried, unused)
//   AbstractGame (abstract_game.h:12-67)                     same eight virtuals
//   VPRNetModel (vprnet.h:26-140)                            one resident engine per model; batch {Initial,Recurrent}Inference,
//                                                             LoadCheckpoint / SaveCheckpoint on the reference's own .pt files
//   DeviceManager / DeviceLoan (device_manager.h:29-109)     least-outstanding-requests loans over the models
//   Evaluator, VPRNetEvaluator (vprnet_evaluator.h:35-102)   blocking one-request calls coalesced into engine batches by the
//                                                             callers themselves (leader/follower; no runner threads or promises)
//   play_game (self_play.cpp:84-183)                         SelfPlayCohort: all live games searched in one lock-step batch per move
//   reanalyze (self_play.cpp:186-225)                        ReanalyzeCohort: position i of every sampled history in one batch
//   algorithm::MCTS (mcts.h:113-187)                          run(): VPRNetEvaluator -> lock-step device search of the model's
//                                                             cohort; any other Evaluator subclass -> device tree kernels with
//                                                             host callbacks per simulation (MZB_EVAL_HOST)
//
// The reference's include paths resolve to this file through include/compat/muzero-cpp/*.h, so tests/mcts_test.cpp of the
// reference compiles UNMODIFIED with -Iinclude/compat (tests/test_compat.py builds and runs it on the GPU).
// MCTS::select_child / ucb_score / backpropogate take hand-built MCTSNode records like the reference's and run them through the
// device step (mzb_debug_*): the arithmetic under test is the tree kernel's, not a host copy.
#pragma once
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <filesystem>
#include <functional>
#include <limits>
#include <map>
#include <numeric>
#include <random>
#include <string>

#include "mzb200_facade.hpp"

namespace muzero_cpp {

namespace types {   // types.h:18-78
using Player = int;
using Action = int;
constexpr Player InvalidPlayer = -1;
constexpr Action InvalidAction = -1;
using Observation = std::vector<float>;
struct ObservationShape {
    int c, h, w;
    bool operator==(const ObservationShape& o) const { return c == o.c && h == o.h && w == o.w; }
    bool operator!=(const ObservationShape& o) const { return !(*this == o); }
};
using ActionRepresentationFunction = std::function<Observation(Action)>;
using SoftmaxTemperatureFunction = std::function<double(int)>;
enum class OpponentTypes { Self, Random, Expert, Human };
enum CheckpointStep { kMostRecentCheckpointStep = -1, kBestPerformanceCheckpointStep = -2 };
struct StepReturn { Observation observation; double reward; bool done; };
struct Batch {   // types.h:81-90: what PrioritizedReplayBuffer::sample hands to Learn
    std::vector<double> priorities;
    std::vector<int> indices;
    std::vector<Action> actions;
    Observation stacked_observations;
    std::vector<double> target_rewards, target_values, target_policies, gradient_scale;
    int num_samples = 0;
};
using MCTSReturn = mzb200::MCTSReturn;   // same five fields in the same order (types.h:70-76)
constexpr double NINF_D = std::numeric_limits<double>::lowest();
constexpr double INF_D = std::numeric_limits<double>::max();
}  // namespace types

namespace muzero_config {   // config.h:15-117, field for field
struct MuZeroNetworkConfig {
    double learning_rate = 3e-4, l2_weight_decay = 1e-4;
    bool downsample = false, normalize_hidden_states = true;
    int resnet_channels = 128, representation_blocks = 4, dynamics_blocks = 4, prediction_blocks = 4;
    int reward_reduced_channels = 16, policy_reduced_channels = 16, value_reduced_channels = 16;
    std::vector<int> reward_head_layers{64}, policy_head_layers{64}, value_head_layers{64};
};
struct MuZeroConfig {
    int seed = 0, checkpoint_interval = 100, model_sync_interval = 100;
    std::string path = "/opt/muzero-cpp/", devices = "cpu:0";
    bool explicit_learning = false;
    int num_actors = 1, num_reanalyze_actors = 0, num_evaluators = 1;
    int initial_inference_batch_size = 1, recurrent_inference_batch_size = 1;
    int initial_inference_threads = 1, recurrent_inference_threads = 1;
    int max_training_steps = 100000;
    bool resume = false;
    types::CheckpointStep testing_checkpoint = types::CheckpointStep::kMostRecentCheckpointStep;
    types::ObservationShape observation_shape{1, 1, 1};
    std::vector<types::Action> action_space;
    int action_channels = 1, num_players = 1, stacked_observations = 0;
    double value_upperbound = types::NINF_D, value_lowerbound = types::INF_D;
    double min_reward = -300, max_reward = 300, min_value = -300, max_value = 300;
    bool use_contractive_mapping = true;
    types::Player muzero_player = 0;
    types::OpponentTypes opponent_type = types::OpponentTypes::Self;
    int max_moves = -1, num_simulations = 50;
    double discount = 0.997, dirichlet_alpha = 0.3, dirichlet_epsilon = 0.25, pb_c_base = 19652, pb_c_init = 1.25;
    int batch_size = 128, min_sample_size = 256;
    double value_loss_weight = 1;
    int td_steps = 10, num_unroll_steps = 5, max_history_len = -1;
    double train_reanalyze_ratio = 0;
    int replay_buffer_size = 100000, reanalyze_buffer_size = 100000;
    double per_alpha = 1, per_beta = 1, per_epsilon = 0.01, per_beta_increment = 0.001;
    MuZeroNetworkConfig network_config;
    types::ActionRepresentationFunction action_representation_initial, action_representation_recurrent;
    types::SoftmaxTemperatureFunction visit_softmax_temperature;
};
}  // namespace muzero_config

class AbstractGame {   // abstract_game.h:12-67
public:
    virtual ~AbstractGame() = default;
    virtual types::Observation reset() = 0;
    virtual types::StepReturn step(types::Action action) = 0;
    virtual types::Player to_play() const = 0;
    virtual std::vector<types::Action> legal_actions() const = 0;
    virtual types::Action expert_action() = 0;
    virtual types::Action human_to_action() = 0;
    virtual void render() {}
    virtual std::string action_to_string(types::Action action) const = 0;
};

namespace detail {
[[noreturn]] inline void fatal(const std::string& msg) {
    std::fprintf(stderr, "mzb200: %s\n", msg.c_str());
    std::abort();
}
inline void copy_layers(const std::vector<int>& src, int32_t* dst, int32_t* n) {
    if ((int)src.size() > MZB_MAX_HEAD_LAYERS) fatal("more head layers than MZB_MAX_HEAD_LAYERS");
    *n = (int32_t)src.size();
    std::copy(src.begin(), src.end(), dst);
}
// MuZeroConfig -> the POD the C ABI takes.  action_planes stays NULL (the examples' constant plane action / num_actions)
// unless action_representation_recurrent is set; the caller owns `planes`.
inline mzb_config to_pod(const muzero_config::MuZeroConfig& c, int max_trees, int precision, int evaluator, std::vector<float>* planes) {
    mzb_config p;
    std::memset(&p, 0, sizeof(p));
    p.observation_c = c.observation_shape.c; p.observation_h = c.observation_shape.h; p.observation_w = c.observation_shape.w;
    p.num_actions = (int32_t)c.action_space.size();
    for (size_t i = 0; i < c.action_space.size(); ++i)
        if (c.action_space[i] != (types::Action)i) fatal("action_space must be {0, ..., num_actions-1} (config.h:55)");
    p.action_channels = c.action_channels; p.num_players = c.num_players; p.stacked_observations = c.stacked_observations;
    p.value_upperbound = c.value_upperbound; p.value_lowerbound = c.value_lowerbound;
    p.min_reward = c.min_reward; p.max_reward = c.max_reward; p.min_value = c.min_value; p.max_value = c.max_value;
    p.use_contractive_mapping = c.use_contractive_mapping;
    p.num_simulations = c.num_simulations; p.discount = c.discount;
    p.dirichlet_alpha = c.dirichlet_alpha; p.dirichlet_epsilon = c.dirichlet_epsilon; p.pb_c_base = c.pb_c_base; p.pb_c_init = c.pb_c_init;
    p.seed = c.seed;
    const muzero_config::MuZeroNetworkConfig& n = c.network_config;
    p.downsample = n.downsample; p.normalize_hidden_states = n.normalize_hidden_states;
    p.resnet_channels = n.resnet_channels; p.representation_blocks = n.representation_blocks;
    p.dynamics_blocks = n.dynamics_blocks; p.prediction_blocks = n.prediction_blocks;
    p.reward_reduced_channels = n.reward_reduced_channels; p.policy_reduced_channels = n.policy_reduced_channels;
    p.value_reduced_channels = n.value_reduced_channels;
    copy_layers(n.reward_head_layers, p.reward_head_layers, &p.n_reward_head_layers);
    copy_layers(n.policy_head_layers, p.policy_head_layers, &p.n_policy_head_layers);
    copy_layers(n.value_head_layers, p.value_head_layers, &p.n_value_head_layers);
    if (c.action_representation_recurrent && planes) {   // vprnet.cpp:143: evaluated once per action instead of once per request
        planes->clear();
        for (types::Action a : c.action_space) {
            const types::Observation plane = c.action_representation_recurrent(a);
            planes->insert(planes->end(), plane.begin(), plane.end());
        }
        p.action_planes = planes->data();
    }
    p.max_trees = max_trees; p.precision = precision; p.evaluator = evaluator;
    return p;
}
// "cuda:3" / "cuda" -> ordinal.  There is no CPU path: the reference's "cpu" / "/cpu:0" device strings fail loudly.
inline int cuda_ordinal(const std::string& device) {
    const size_t at = device.find("cuda");
    if (at == std::string::npos) fatal("device \"" + device + "\": libmzb200 has no CPU path, use cuda:N");
    const size_t colon = device.find(':', at);
    return colon == std::string::npos ? 0 : std::atoi(device.c_str() + colon + 1);
}
}  // namespace detail

// vprnet.h:26-140.  One model = one GPU-resident engine: weights, node pool and max_trees tree slots, where max_trees covers
// every MCTS object the reference would create for this config (actors + reanalyze actors + evaluators, self_play.cpp:228-275)
// and both inference batch sizes.
class VPRNetModel {
public:
    struct InitialInferenceInputs { types::Observation stacked_observation; };               // vprnet.h:52-55
    struct RecurrentInferenceInputs { types::Action action; types::Observation encoded_state; };   // vprnet.h:58-61
    struct InferenceOutputs {                                                                  // vprnet.h:64-69
        double value; double reward; std::vector<double> policy; types::Observation encoded_state;
    };

    // precision: MZB_BF16 (tensor-core towers) or MZB_FP32_CHECK; MZB200_PRECISION=fp32 in the environment overrides.
    VPRNetModel(const muzero_config::MuZeroConfig& config, const std::string& device = "cuda:0", int precision = MZB_BF16)
        : st_(new State) {
        const char* env = std::getenv("MZB200_PRECISION");
        if (env && std::string(env) == "fp32") precision = MZB_FP32_CHECK;
        st_->device = device;
        st_->path = config.path + "/checkpoints/";   // vprnet.cpp:32
        st_->num_unroll_steps = config.num_unroll_steps; st_->value_loss_weight = config.value_loss_weight;
        st_->learning_rate = config.network_config.learning_rate; st_->l2_weight_decay = config.network_config.l2_weight_decay;
        const int trees = std::max({config.num_actors + config.num_reanalyze_actors + config.num_evaluators,
                                    config.initial_inference_batch_size, config.recurrent_inference_batch_size, 1});
        const mzb_config pod = detail::to_pod(config, trees, precision, MZB_EVAL_NETWORK, &st_->planes);
        st_->engine = std::make_shared<mzb200::Engine>(pod, detail::cuda_ordinal(device));
        const char* dl = std::getenv("MZB200_COHORT_DEADLINE_US");
        st_->cohort = std::make_shared<mzb200::Cohort>(st_->engine, std::chrono::microseconds(dl ? std::atol(dl) : 1000));
        st_->shapes = st_->engine->shapes();
        st_->num_actions = pod.num_actions;
    }
    VPRNetModel(VPRNetModel&&) = default;
    VPRNetModel& operator=(VPRNetModel&&) = default;
    VPRNetModel(const VPRNetModel&) = delete;
    VPRNetModel& operator=(const VPRNetModel&) = delete;

    void print() const {
        std::printf("mzb200 engine on %s: hidden %dx%dx%d, weights %lld B, node pool %lld B\n", st_->device.c_str(), st_->shapes.enc_c,
                    st_->shapes.enc_h, st_->shapes.enc_w, (long long)st_->shapes.weight_blob_bytes, (long long)st_->shapes.node_pool_bytes);
    }

    std::vector<InferenceOutputs> InitialInference(std::vector<InitialInferenceInputs>& inputs) {   // vprnet.cpp:77-129
        const int B = (int)inputs.size(), flat = st_->shapes.initial_flat_size;
        std::vector<float> rows((size_t)B * flat);
        for (int b = 0; b < B; ++b) {
            if ((int)inputs[b].stacked_observation.size() != flat) detail::fatal("InitialInference: observation of the wrong size");
            std::copy(inputs[b].stacked_observation.begin(), inputs[b].stacked_observation.end(), rows.begin() + (size_t)b * flat);
        }
        Outputs o(B, st_->num_actions, st_->shapes.recurrent_flat_size);
        {
            std::lock_guard<std::mutex> call(st_->engine->call_mutex());
            mzb200::check(mzb_initial_inference(st_->engine->handle(), rows.data(), B, o.value.data(), o.reward.data(), o.policy.data(), o.hidden.data()),
                          "mzb_initial_inference");
        }
        return o.unpack();
    }
    std::vector<InferenceOutputs> RecurrentInference(std::vector<RecurrentInferenceInputs>& inputs) {   // vprnet.cpp:131-189
        const int B = (int)inputs.size(), flat = st_->shapes.recurrent_flat_size;
        std::vector<float> rows((size_t)B * flat);
        std::vector<int32_t> actions(B);
        for (int b = 0; b < B; ++b) {
            if ((int)inputs[b].encoded_state.size() != flat) detail::fatal("RecurrentInference: encoded state of the wrong size");
            std::copy(inputs[b].encoded_state.begin(), inputs[b].encoded_state.end(), rows.begin() + (size_t)b * flat);
            actions[b] = inputs[b].action;
        }
        Outputs o(B, st_->num_actions, flat);
        {
            std::lock_guard<std::mutex> call(st_->engine->call_mutex());
            mzb200::check(mzb_recurrent_inference(st_->engine->handle(), rows.data(), actions.data(), B, o.value.data(), o.reward.data(),
                                                  o.policy.data(), o.hidden.data()), "mzb_recurrent_inference");
        }
        return o.unpack();
    }

    // vprnet.cpp:66-74: reads "<path>.pt" as torch::save wrote it (vprnet.cpp:58-63) with the library's own archive reader — no
    // libtorch.  The "-optimizer.pt" file next to it is the learner's and is not read.
    void LoadCheckpoint(const std::string& path) {
        std::lock_guard<std::mutex> call(st_->engine->call_mutex());
        mzb200::check(mzb_load_checkpoint(st_->engine->handle(), (path + ".pt").c_str()), "mzb_load_checkpoint");
    }
    void LoadCheckpoint(int step) { LoadCheckpoint(st_->path + "checkpoint-" + std::to_string(step)); }
    // vprnet.cpp:58-63: writes "<config.path>/checkpoints/checkpoint-<step>.pt" in the format torch::save gives the module (so
    // that a reference process, --resume or --test can torch::load it) and returns the path without the extension, like the
    // reference.  What is written: the tensors this model was last loaded with (LoadCheckpoint / LoadWeights).  The
    // "-optimizer.pt" file is the learner's (Adam state) and is not written here.
    std::string SaveCheckpoint(int step) {
        const std::string full_path = st_->path + "checkpoint-" + std::to_string(step);
        std::error_code ec;
        std::filesystem::create_directories(st_->path, ec);   // the reference's constructor creates the directory (vprnet.cpp:33)
        if (ec) detail::fatal("SaveCheckpoint: cannot create " + st_->path + ": " + ec.message());
        std::lock_guard<std::mutex> call(st_->engine->call_mutex());
        mzb200::check(mzb_save_checkpoint(st_->engine->handle(), (full_path + ".pt").c_str()), "mzb_save_checkpoint");
        return full_path;
    }
    // The same hand-over without a file: the module's named fp32 tensors (a learner in the same process, tests).
    void LoadWeights(const std::vector<mzb_tensor_desc>& tensors) {
        std::lock_guard<std::mutex> call(st_->engine->call_mutex());
        st_->engine->LoadWeights(tensors);
    }
    // vprnet.cpp:192-311, the forward half (first slice): the K + 1 unrolled TRAIN-mode network calls and the
    // loss of Learn(batch) on the GPU, returned as the reference's LossInfo.  Named LearnForward because it stops there: no
    // backward pass, no Adam step — the weights stay as they are (BatchNorm running statistics of the training state move).
    struct LossInfo {   // vprnet.h:33-50
        double total_loss = 0, value = 0, policy = 0, reward = 0;
        std::vector<int> indices;
        std::vector<double> errors;
    };
    template <class Batch>   // types::Batch (types.h:81-90) or any struct with its fields
    LossInfo LearnForward(const Batch& inputs, int num_unroll_steps, double value_loss_weight) {
        std::vector<int32_t> actions(inputs.actions.begin(), inputs.actions.end());
        LossInfo li;
        li.indices = inputs.indices;
        li.errors.assign(inputs.num_samples, 0.0);
        mzb_batch b{};
        b.num_samples = inputs.num_samples; b.actions = actions.data(); b.stacked_observations = inputs.stacked_observations.data();
        b.target_values = inputs.target_values.data(); b.target_rewards = inputs.target_rewards.data();
        b.target_policies = inputs.target_policies.data(); b.gradient_scale = inputs.gradient_scale.data(); b.priorities = inputs.priorities.data();
        mzb_loss_info out{};
        out.errors = li.errors.data();
        std::lock_guard<std::mutex> call(st_->engine->call_mutex());
        mzb200::check(mzb_learn_forward(st_->engine->handle(), &b, num_unroll_steps, value_loss_weight, &out), "mzb_learn_forward");
        li.total_loss = out.total_loss; li.value = out.value; li.policy = out.policy; li.reward = out.reward;
        return li;
    }
    // vprnet.cpp:192-311, all of it: forward, loss, backward with Learn's gradient hooks, one torch::optim::Adam step with
    // config.network_config.learning_rate / l2_weight_decay (vprnet.cpp:35-36), fp32 on the GPU (mzb_learn_step).  Like the
    // reference's model, this one answers inference with the trained tensors right after the call (`commit`, the default); a learner
    // that only hands weights over every checkpoint_interval steps passes commit = false and calls CommitLearnedWeights() then.
    template <class Batch>
    LossInfo Learn(Batch& inputs, bool commit = true) {
        std::vector<int32_t> actions(inputs.actions.begin(), inputs.actions.end());
        LossInfo li;
        li.indices = inputs.indices;
        li.errors.assign(inputs.num_samples, 0.0);
        mzb_batch b{};
        b.num_samples = inputs.num_samples; b.actions = actions.data(); b.stacked_observations = inputs.stacked_observations.data();
        b.target_values = inputs.target_values.data(); b.target_rewards = inputs.target_rewards.data();
        b.target_policies = inputs.target_policies.data(); b.gradient_scale = inputs.gradient_scale.data(); b.priorities = inputs.priorities.data();
        mzb_loss_info out{};
        out.errors = li.errors.data();
        std::lock_guard<std::mutex> call(st_->engine->call_mutex());
        mzb200::check(mzb_learn_step(st_->engine->handle(), &b, st_->num_unroll_steps, st_->value_loss_weight, st_->learning_rate, st_->l2_weight_decay, &out),
                      "mzb_learn_step");
        if (commit) mzb200::check(mzb_learn_commit(st_->engine->handle()), "mzb_learn_commit");
        li.total_loss = out.total_loss; li.value = out.value; li.policy = out.policy; li.reward = out.reward;
        return li;
    }
    void CommitLearnedWeights() {
        std::lock_guard<std::mutex> call(st_->engine->call_mutex());
        mzb200::check(mzb_learn_commit(st_->engine->handle()), "mzb_learn_commit");
    }
    const std::string Device() const { return st_->device; }

    // mzb200 side
    std::shared_ptr<mzb200::Engine> engine() const { return st_->engine; }
    std::shared_ptr<mzb200::Cohort> cohort() const { return st_->cohort; }

private:
    struct Outputs {
        int B, A, S;
        std::vector<double> value, reward, policy;
        std::vector<float> hidden;
        Outputs(int b, int a, int s) : B(b), A(a), S(s), value(b), reward(b), policy((size_t)b * a), hidden((size_t)b * s) {}
        std::vector<InferenceOutputs> unpack() const {
            std::vector<InferenceOutputs> out(B);
            for (int b = 0; b < B; ++b) {
                out[b].value = value[b]; out[b].reward = reward[b];
                out[b].policy.assign(policy.begin() + (size_t)b * A, policy.begin() + (size_t)(b + 1) * A);
                out[b].encoded_state.assign(hidden.begin() + (size_t)b * S, hidden.begin() + (size_t)(b + 1) * S);
            }
            return out;
        }
    };
    struct State {
        std::string device, path;
        std::vector<float> planes;
        std::shared_ptr<mzb200::Engine> engine;
        std::shared_ptr<mzb200::Cohort> cohort;
        mzb_shapes shapes;
        int num_actions = 0;
        int num_unroll_steps = 5;
        double value_loss_weight = 1, learning_rate = 3e-4, l2_weight_decay = 1e-4;
    };
    std::unique_ptr<State> st_;
};

// device_manager.h:29-109: hands out the model with the fewest outstanding requests; device 0 is left to the learner while
// SetLearning(true) and more than one device exists.
class DeviceManager {
public:
    DeviceManager() = default;
    void AddDevice(VPRNetModel model) { slots_.emplace_back(new Slot{std::move(model), 0, 0}); }   // not thread safe (as the reference)

    class DeviceLoan {
    public:
        DeviceLoan(DeviceLoan&& o) noexcept : owner_(o.owner_), model_(o.model_), id_(o.id_), n_(o.n_) { o.owner_ = nullptr; }
        DeviceLoan& operator=(DeviceLoan&& o) noexcept {
            if (this != &o) { release(); owner_ = o.owner_; model_ = o.model_; id_ = o.id_; n_ = o.n_; o.owner_ = nullptr; }
            return *this;
        }
        DeviceLoan(const DeviceLoan&) = delete;
        DeviceLoan& operator=(const DeviceLoan&) = delete;
        ~DeviceLoan() { release(); }
        VPRNetModel* operator->() { return model_; }
        int device_id() const { return id_; }

    private:
        friend class DeviceManager;
        DeviceLoan(DeviceManager* owner, VPRNetModel* model, int id, int n) : owner_(owner), model_(model), id_(id), n_(n) {}
        void release() { if (owner_) owner_->give_back(id_, n_); owner_ = nullptr; }
        DeviceManager* owner_;
        VPRNetModel* model_;
        int id_, n_;
    };

    DeviceLoan Get(int requests, int device_id = -1) {
        std::lock_guard<std::mutex> lk(mu_);
        if (slots_.empty()) detail::fatal("DeviceManager::Get with no devices");
        if (device_id < 0) {
            const int first = (learning_ && slots_.size() > 1) ? 1 : 0;
            device_id = first;
            for (int i = first + 1; i < (int)slots_.size(); ++i)
                if (slots_[i]->outstanding < slots_[device_id]->outstanding) device_id = i;
        }
        slots_[device_id]->outstanding += requests;
        return DeviceLoan(this, &slots_[device_id]->model, device_id, requests);
    }
    void SetLearning(bool value) { std::lock_guard<std::mutex> lk(mu_); learning_ = value; }
    int Count() const { return (int)slots_.size(); }

    // mzb200 side: an MCTS object lives on ONE device for its lifetime (its tree slot and RNG stream are resident there);
    // it goes to the device with the fewest resident searchers, honouring the learner's reservation of device 0.
    VPRNetModel* attach_searcher(int* device_id) {
        std::lock_guard<std::mutex> lk(mu_);
        if (slots_.empty()) detail::fatal("DeviceManager has no devices");
        const int first = (learning_ && slots_.size() > 1) ? 1 : 0;
        int best = first;
        for (int i = first + 1; i < (int)slots_.size(); ++i)
            if (slots_[i]->searchers < slots_[best]->searchers) best = i;
        ++slots_[best]->searchers;
        *device_id = best;
        return &slots_[best]->model;
    }
    void detach_searcher(int device_id) { std::lock_guard<std::mutex> lk(mu_); --slots_[device_id]->searchers; }

private:
    struct Slot { VPRNetModel model; int outstanding; int searchers; };
    void give_back(int id, int n) { std::lock_guard<std::mutex> lk(mu_); slots_[id]->outstanding -= n; }
    std::vector<std::unique_ptr<Slot>> slots_;   // stable addresses: loans and searchers keep VPRNetModel*
    bool learning_ = false;
    std::mutex mu_;
};

class Evaluator {   // vprnet_evaluator.h:35-41
public:
    virtual ~Evaluator() = default;
    virtual VPRNetModel::InferenceOutputs InitialInference(const types::Observation& stacked_observations) = 0;
    virtual VPRNetModel::InferenceOutputs RecurrentInference(const types::Observation& hidden_state, types::Action action) = 0;
};

// vprnet_evaluator.h:44-102, vprnet_evaluator.cpp:25-178.  Same contract — many threads call the one-request methods and block
// until their answer is ready; requests are served in batches of up to batch_size, or fewer once the oldest has waited 1 ms —
// but no runner threads, queue items or promises: the caller that completes a batch (or sees the deadline pass) becomes its
// leader, borrows the least-loaded model, runs ONE engine call for the batch and wakes the others.  `threads` bounds how many
// batches may be in flight at once, which is what the reference's runner-thread counts control.
class VPRNetEvaluator : public Evaluator {
public:
    explicit VPRNetEvaluator(DeviceManager* device_manager, int batch_size_initial, int initial_inference_threads,
                             int batch_size_recurrent, int recurrent_inference_threads)
        : dm_(*device_manager), initial_(batch_size_initial, initial_inference_threads), recurrent_(batch_size_recurrent, recurrent_inference_threads) {}

    VPRNetModel::InferenceOutputs InitialInference(const types::Observation& stacked_observations) override {
        Request rq;
        rq.input = &stacked_observations;
        return initial_.submit(rq, [this](std::vector<Request*>& batch) {
            std::vector<VPRNetModel::InitialInferenceInputs> in(batch.size());
            for (size_t i = 0; i < batch.size(); ++i) in[i].stacked_observation = *batch[i]->input;
            auto out = dm_.Get((int)batch.size())->InitialInference(in);
            for (size_t i = 0; i < batch.size(); ++i) batch[i]->output = std::move(out[i]);
        });
    }
    VPRNetModel::InferenceOutputs RecurrentInference(const types::Observation& hidden_state, types::Action action) override {
        Request rq;
        rq.input = &hidden_state; rq.action = action;
        return recurrent_.submit(rq, [this](std::vector<Request*>& batch) {
            std::vector<VPRNetModel::RecurrentInferenceInputs> in(batch.size());
            for (size_t i = 0; i < batch.size(); ++i) { in[i].action = batch[i]->action; in[i].encoded_state = *batch[i]->input; }
            auto out = dm_.Get((int)batch.size())->RecurrentInference(in);
            for (size_t i = 0; i < batch.size(); ++i) batch[i]->output = std::move(out[i]);
        });
    }
    DeviceManager& device_manager() { return dm_; }
    // batches served so far / requests served so far (the reference logs neither; used by tests to see that batching happens)
    long batches_served() const { return initial_.batches + recurrent_.batches; }
    long requests_served() const { return initial_.requests + recurrent_.requests; }

private:
    struct Request {
        const types::Observation* input = nullptr;
        types::Action action = 0;
        VPRNetModel::InferenceOutputs output;
        bool taken = false, done = false;
    };
    struct BatchPoint {
        BatchPoint(int batch_size, int threads) : batch_size(std::max(1, batch_size)), max_in_flight(std::max(1, threads)) {}
        template <class Run>
        VPRNetModel::InferenceOutputs submit(Request& rq, Run run) {
            std::unique_lock<std::mutex> lk(mu);
            if (pending.empty()) oldest = std::chrono::steady_clock::now();
            pending.push_back(&rq);
            cv.notify_all();
            while (!rq.done) {
                const bool due = (int)pending.size() >= batch_size || std::chrono::steady_clock::now() - oldest >= deadline;
                if (!rq.taken && due && in_flight < max_in_flight) {
                    std::vector<Request*> batch;
                    const size_t n = std::min(pending.size(), (size_t)batch_size);
                    batch.assign(pending.begin(), pending.begin() + n);
                    pending.erase(pending.begin(), pending.begin() + n);
                    for (Request* r : batch) r->taken = true;
                    if (!pending.empty()) oldest = std::chrono::steady_clock::now();
                    ++in_flight;
                    lk.unlock();
                    run(batch);
                    lk.lock();
                    --in_flight; ++batches; requests += (long)batch.size();
                    for (Request* r : batch) r->done = true;
                    cv.notify_all();
                } else if (!rq.taken && in_flight < max_in_flight) {
                    cv.wait_until(lk, oldest + deadline);
                } else {
                    cv.wait(lk);
                }
            }
            return std::move(rq.output);
        }
        const int batch_size, max_in_flight;
        const std::chrono::microseconds deadline{1000};   // vprnet_evaluator.cpp:118
        std::mutex mu;
        std::condition_variable cv;
        std::vector<Request*> pending;
        std::chrono::steady_clock::time_point oldest;
        int in_flight = 0;
        long batches = 0, requests = 0;
    };
    DeviceManager& dm_;
    BatchPoint initial_, recurrent_;
};

namespace util {
// util.cpp:28-43: N draws of Gamma(alpha, 1) from the caller's generator, normalised.  The arithmetic is libstdc++'s
// std::gamma_distribution, as in the reference — the engine's own root noise (engine_search.cu) draws the same way on the tree
// slot's generator state.
inline std::vector<double> sample_dirichlet(double alpha, int num_samples, std::mt19937& rng) {
    std::vector<double> samples;
    samples.reserve(num_samples);
    std::gamma_distribution<double> gamma_dist(alpha, 1.0);
    for (int i = 0; i < num_samples; ++i) samples.push_back(gamma_dist(rng));
    const double sum = std::reduce(samples.begin(), samples.end());
    for (double& x : samples) x /= sum;
    return samples;
}
}  // namespace util

namespace algorithm {

// mcts.h:79-112 (mcts.cpp:205-220): the running min/max a search normalises Q values with.  Inside MCTS::run the statistics live
// in the tree slot on the device (tree.cu: updated by the backup half, read by pUCT); this is the same record for callers that
// hold one themselves, as tests/mcts_test.cpp:156-188 does.
class MinMaxStats {
public:
    MinMaxStats(double minimum, double maximum) : init_minimum_(minimum), init_maximum_(maximum), minimum_(minimum), maximum_(maximum) {}
    MinMaxStats() : MinMaxStats(types::INF_D, types::NINF_D) {}
    void reset() { maximum_ = init_maximum_; minimum_ = init_minimum_; }
    void update(double value) { maximum_ = std::max(maximum_, value); minimum_ = std::min(minimum_, value); }
    double normalize(double value) const { return (maximum_ > minimum_) ? (value - minimum_) / (maximum_ - minimum_) : value; }
    double minimum() const { return minimum_; }
    double maximum() const { return maximum_; }

private:
    double init_minimum_, init_maximum_, minimum_, maximum_;
};

// mcts.h:19-75 as a HOST record: what a caller (tests/mcts_test.cpp) assembles by hand and passes to the three public helpers
// of MCTS below.  The search itself never builds these — its trees are structure-of-arrays pools in HBM — so only the
// bookkeeping members exist here; add_exploration_noise and get_search_statistics run inside MCTS::run (tree_root_kernel,
// tree_stats_kernel) and are pinned through its goldens.
struct MCTSNode {
    int visit_count = 0;
    types::Player to_play = types::InvalidPlayer;
    types::Action action = types::InvalidAction;
    double prior = 0, value_sum = 0, reward = 0;
    std::vector<MCTSNode> children;
    types::Observation encoded_state;
    MCTSNode(types::Action action, double prior) : action(action), prior(prior) {}
    MCTSNode() = delete;
    bool is_expanded() const { return !children.empty(); }                              // mcts.cpp:146-148
    double value() const { return visit_count == 0 ? 0 : value_sum / visit_count; }     // mcts.cpp:151-153
    void expand(const std::vector<types::Action>& actions, types::Player player, double r, const std::vector<double>& policy_probs,
                const types::Observation& state) {                                      // mcts.cpp:157-167
        to_play = player; reward = r; encoded_state = state;
        children.reserve(actions.size());
        for (types::Action a : actions) children.emplace_back(a, policy_probs[a]);
    }
    // mcts.cpp:170-176 on the caller's generator (inside MCTS::run: tree_root_kernel mixes the host-drawn noise in)
    void add_exploration_noise(double dirichlet_alpha, double dirichlet_epsilon, std::mt19937& rng) {
        const std::vector<double> noise = util::sample_dirichlet(dirichlet_alpha, (int)children.size(), rng);
        for (size_t i = 0; i < children.size(); ++i) children[i].prior = children[i].prior * (1 - dirichlet_epsilon) + (noise[i] * dirichlet_epsilon);
    }
    // mcts.cpp:179-200 (inside MCTS::run: tree_stats_kernel)
    void get_search_statistics(const std::vector<types::Action>& action_space, std::vector<types::Action>& child_actions,
                               std::vector<double>& children_relative_visit, std::vector<double>& emperical_policy) const {
        int sum_visits = 0;
        emperical_policy = std::vector<double>(action_space.size(), 0);
        for (const MCTSNode& c : children) {
            child_actions.push_back(c.action);
            children_relative_visit.push_back(c.visit_count);
            emperical_policy[c.action] = c.visit_count;
            sum_visits += c.visit_count;
        }
        for (double& v : children_relative_visit) v /= sum_visits;
        for (double& v : emperical_policy) v /= sum_visits;
    }
};

// mcts.h:113-187.  Thread-confined, one per actor, as in the reference.
class MCTS {
public:
    MCTS(const muzero_config::MuZeroConfig& config, int seed, std::shared_ptr<Evaluator> vpr_eval)
        : eval_(std::move(vpr_eval)), players_(config.num_players), mm_min_(config.value_lowerbound), mm_max_(config.value_upperbound) {
        if (auto net = std::dynamic_pointer_cast<VPRNetEvaluator>(eval_)) {
            // the search runs where the weights are: a tree slot of the chosen model's cohort
            dm_ = &net->device_manager();
            VPRNetModel* model = dm_->attach_searcher(&device_id_);
            device_search_.reset(new mzb200::MCTS(model->cohort(), seed));
            dm_engine_ = model->engine();
        } else {
            // a user-defined Evaluator (tests/mcts_test.cpp:16-27): a private one-tree engine whose two inference calls come back here
            const mzb_config pod = detail::to_pod(config, 1, MZB_FP32_CHECK, MZB_EVAL_HOST, nullptr);
            host_engine_.reset(new mzb200::Engine(pod, 0));
            node_state_.resize(config.num_simulations + 1);
            mzb_host_evaluator cb;
            cb.user = this; cb.initial = &MCTS::host_initial; cb.recurrent = &MCTS::host_recurrent;
            mzb200::check(mzb_set_host_evaluator(host_engine_->handle(), &cb), "mzb_set_host_evaluator");
            const int32_t sd = seed;
            mzb200::check(mzb_search_seed(host_engine_->handle(), 0, 1, &sd), "mzb_search_seed");
            num_actions_ = pod.num_actions;
        }
    }
    MCTS() = delete;
    MCTS(const MCTS&) = delete;
    MCTS& operator=(const MCTS&) = delete;
    ~MCTS() {
        device_search_.reset();
        if (dm_) dm_->detach_searcher(device_id_);
    }

    types::MCTSReturn run(const types::Observation& stacked_observation, const std::vector<types::Action>& legal_actions,
                          types::Player to_play, bool add_exploration_noise) {
        if (device_search_) return device_search_->run(stacked_observation, legal_actions, to_play, add_exploration_noise);
        const int A = num_actions_;
        const uint32_t mask = mzb200::legal_mask_of(legal_actions, A);
        const int32_t tp = to_play;
        const uint8_t noise = add_exploration_noise ? 1 : 0;
        // the engine's row is initial_flat_size floats; a test evaluator may be handed any observation, so pad / trim a copy
        const mzb_shapes sh = host_engine_->shapes();
        std::vector<float> row(sh.initial_flat_size, 0.f);
        std::copy(stacked_observation.begin(), stacked_observation.begin() + std::min(row.size(), stacked_observation.size()), row.begin());
        root_obs_ = &stacked_observation;
        int32_t n_children = 0, depth = 0;
        double root = 0;
        std::vector<int32_t> actions(A);
        std::vector<double> rel(A), pol(A);
        mzb_search_result out{};
        out.root_value = &root; out.max_tree_depth = &depth; out.n_children = &n_children;
        out.child_actions = actions.data(); out.children_relative_visit = rel.data(); out.emperical_policy = pol.data();
        mzb200::check(mzb_search_batch(host_engine_->handle(), row.data(), &mask, &tp, &noise, 1, &out), "mzb_search_batch");
        types::MCTSReturn r;
        r.root_value = root; r.max_tree_depth = depth; r.emperical_policy = pol;
        r.children_relative_visit.assign(rel.begin(), rel.begin() + n_children);
        r.child_actions.assign(actions.begin(), actions.begin() + n_children);
        return r;
    }

    // ---- the three public helpers (mcts.h:149-172), run by the DEVICE step on an upload of the caller's nodes ----
    // select_child (mcts.cpp:110-124): shuffles node->children with this object's generator, returns the first strict pUCT maximum
    MCTSNode* select_child(MCTSNode* node) {
        std::vector<int32_t> tags;
        const int chosen = device_select(node, /*restore_rng*/ 0, &tags, nullptr);
        std::vector<MCTSNode> shuffled;
        shuffled.reserve(node->children.size());
        for (int32_t tag : tags) shuffled.push_back(std::move(node->children[tag]));
        node->children = std::move(shuffled);
        return &node->children[chosen];
    }
    // ucb_score (mcts.cpp:127-141)
    double ucb_score(MCTSNode* parent, MCTSNode* child) {
        std::vector<int32_t> tags;
        std::vector<double> scores;
        device_select(parent, /*restore_rng*/ 1, &tags, &scores);
        const int32_t want = (int32_t)(child - parent->children.data());
        for (size_t j = 0; j < tags.size(); ++j) if (tags[j] == want) return scores[j];
        detail::fatal("ucb_score: child is not a child of parent");
    }
    // backpropogate (mcts.cpp:90-107); updates this object's MinMaxStats like the reference's member
    void backpropogate(std::vector<MCTSNode*>& search_path, double value, types::Player to_play) {
        const int len = (int)search_path.size();
        const types::Player root_tp = (players_ == 2) ? ((to_play + (len - 1)) % 2) : 0;   // players alternate down the path
        std::vector<int32_t> meta(4 * len), path(len);
        std::vector<double> prior(len), vsum(len), rew(len);
        for (int i = 0; i < len; ++i) {
            const MCTSNode& nd = *search_path[i];
            if (nd.to_play != types::InvalidPlayer && nd.to_play != (root_tp + i) % players_)
                detail::fatal("backpropogate: players do not alternate along the path (the device tree derives to_play from depth)");
            meta[4 * i] = nd.visit_count; meta[4 * i + 1] = nd.action; meta[4 * i + 2] = -1; meta[4 * i + 3] = i;
            prior[i] = nd.prior; vsum[i] = nd.value_sum; rew[i] = nd.reward; path[i] = i;
        }
        mzb_debug_tree t{};
        t.n_nodes = len; t.meta = meta.data(); t.prior = prior.data(); t.value_sum = vsum.data(); t.reward = rew.data();
        t.minmax_minimum = mm_min_; t.minmax_maximum = mm_max_; t.root_to_play = root_tp; t.n_root_children = 0;
        const std::vector<double> policy(helper_actions(), 1.0 / helper_actions());
        double mm[2];
        with_helper_engine([&](mzb_engine* e, int slot) {
            mzb200::check(mzb_debug_tree_write(e, slot, &t), "mzb_debug_tree_write");
            mzb200::check(mzb_debug_backpropagate(e, slot, path.data(), len, value, search_path.back()->reward, policy.data()), "mzb_debug_backpropagate");
            mzb200::check(mzb_debug_tree_read(e, slot, len, meta.data(), nullptr, vsum.data(), nullptr, mm), "mzb_debug_tree_read");
        });
        for (int i = 0; i < len; ++i) { search_path[i]->visit_count = meta[4 * i]; search_path[i]->value_sum = vsum[i]; }
        mm_min_ = mm[0]; mm_max_ = mm[1];
    }

private:
    int helper_actions() const { return host_engine_ ? num_actions_ : dm_engine_->config().num_actions; }
    template <class Fn>
    void with_helper_engine(Fn fn) {
        if (host_engine_) { fn(host_engine_->handle(), 0); return; }
        std::lock_guard<std::mutex> call(dm_engine_->call_mutex());
        fn(dm_engine_->handle(), device_search_->slot());
    }
    // uploads `node` and its children (as leaves), runs the device selection step; returns the chosen child's position and the
    // children's tags (original positions) in post-shuffle order
    int device_select(MCTSNode* node, int restore_rng, std::vector<int32_t>* tags, std::vector<double>* scores) {
        const int n = (int)node->children.size();
        if (n < 1 || n > helper_actions()) detail::fatal("select_child: node has no children (or more than num_actions)");
        std::vector<int32_t> meta(4 * (n + 1));
        std::vector<double> prior(n + 1), vsum(n + 1), rew(n + 1), sc(helper_actions());
        meta[0] = node->visit_count; meta[1] = node->action; meta[2] = 1; meta[3] = 0;
        prior[0] = node->prior; vsum[0] = node->value_sum; rew[0] = node->reward;
        for (int j = 0; j < n; ++j) {
            const MCTSNode& c = node->children[j];
            meta[4 * (j + 1)] = c.visit_count; meta[4 * (j + 1) + 1] = c.action; meta[4 * (j + 1) + 2] = -1; meta[4 * (j + 1) + 3] = j;
            prior[j + 1] = c.prior; vsum[j + 1] = c.value_sum; rew[j + 1] = c.reward;
        }
        mzb_debug_tree t{};
        t.n_nodes = n + 1; t.meta = meta.data(); t.prior = prior.data(); t.value_sum = vsum.data(); t.reward = rew.data();
        t.minmax_minimum = mm_min_; t.minmax_maximum = mm_max_;
        t.root_to_play = (node->to_play >= 0 && node->to_play < players_) ? node->to_play : 0; t.n_root_children = n;
        int32_t chosen = 0;
        with_helper_engine([&](mzb_engine* e, int slot) {
            mzb200::check(mzb_debug_tree_write(e, slot, &t), "mzb_debug_tree_write");
            mzb200::check(mzb_debug_select_child(e, slot, restore_rng, &chosen, sc.data()), "mzb_debug_select_child");
            mzb200::check(mzb_debug_tree_read(e, slot, n + 1, meta.data(), nullptr, nullptr, nullptr, nullptr), "mzb_debug_tree_read");
        });
        tags->resize(n);
        for (int j = 0; j < n; ++j) (*tags)[j] = meta[4 * (j + 1) + 3];
        if (scores) scores->assign(sc.begin(), sc.begin() + n);
        return chosen - 1;
    }
    int store(const VPRNetModel::InferenceOutputs& o, int node, double* value, double* reward, double* policy) {
        if ((int)o.policy.size() != num_actions_) return -1;
        *value = o.value; *reward = o.reward;
        std::copy(o.policy.begin(), o.policy.end(), policy);
        node_state_[node] = o.encoded_state;
        return 0;
    }
    static int host_initial(void* user, int, const int32_t*, const float*, int, double* value, double* reward, double* policy) {
        MCTS* self = static_cast<MCTS*>(user);
        return self->store(self->eval_->InitialInference(*self->root_obs_), 0, value, reward, policy);
    }
    static int host_recurrent(void* user, int, const int32_t*, const int32_t* parent_node, const int32_t* action, int child_node,
                              double* value, double* reward, double* policy) {
        MCTS* self = static_cast<MCTS*>(user);
        return self->store(self->eval_->RecurrentInference(self->node_state_[parent_node[0]], action[0]), child_node, value, reward, policy);
    }

    std::shared_ptr<Evaluator> eval_;
    // VPRNetEvaluator path
    DeviceManager* dm_ = nullptr;
    int device_id_ = -1;
    std::unique_ptr<mzb200::MCTS> device_search_;
    std::shared_ptr<mzb200::Engine> dm_engine_;
    // user-evaluator path
    std::unique_ptr<mzb200::Engine> host_engine_;
    std::vector<types::Observation> node_state_;   // encoded_state of every expanded node, by expansion order
    const types::Observation* root_obs_ = nullptr;
    int num_actions_ = 0;
    // helpers: MinMaxStats(value_lowerbound, value_upperbound) as the reference's member (mcts.cpp:25), advanced by backpropogate
    int players_;
    double mm_min_, mm_max_;
};

}  // namespace algorithm

// ---------------------------------------------------------------------------------------------------------------------
// The callers' side for cohorts of games: play_game (self_play.cpp:84-183) for MANY games at once.
// The reference runs one actor thread per game, each blocking on its own MCTS::run; here every live game of the cohort is
// searched in ONE lock-step device search per move, fed from the device-resident observation ring (one raw frame per game and
// move crosses PCIe instead of the stacked observation), and the moves are drawn by the engine's per-actor generators
// (std::mt19937(config.seed + actor_num), self_play.cpp:233).  Per game, the result is what play_game hands to the
// trajectory queue: the same GameHistory fields, filled in the same order.
namespace types {
struct GameHistory {   // types.h:110-118, the fields play_game fills
    std::vector<Observation> observation_history;
    std::vector<Action> action_history;
    std::vector<double> reward_history;
    std::vector<Player> to_play_history;
    std::vector<std::vector<Action>> legal_actions;
    std::vector<double> root_values;
    std::vector<std::vector<double>> child_visits;
    void store_search_statistics(double root_value, const std::vector<double>& relative_visits) {   // types.cpp:10-13
        root_values.push_back(root_value);
        child_visits.push_back(relative_visits);
    }
};
}  // namespace types

namespace buffer {
// PrioritizedReplayBuffer (replay_buffer.h:212-300) on the engine's GPU: same constructor arguments plus the engine, same methods,
// same batches for the same std::mt19937 state (mzb_replay_*, csrc/replay.cu: sum tree on the host, histories in HBM, one kernel
// launch per batch).  save() / load() (libnop files) are not mirrored.  A host copy of each stored history is kept for sample_game,
// which hands whole histories to reanalyze (self_play.cpp:198-200); it is dropped when the buffer evicts the history.
class PrioritizedReplayBuffer {
public:
    PrioritizedReplayBuffer(const muzero_config::MuZeroConfig& config, int max_size, const std::string& name, std::shared_ptr<mzb200::Engine> engine)
        : engine_(std::move(engine)), name_(name), A_((int)config.action_space.size()) {
        std::vector<float> planes;
        if (config.action_representation_initial)
            for (types::Action a : config.action_space) {
                const types::Observation pl = config.action_representation_initial(a);
                planes.insert(planes.end(), pl.begin(), pl.end());
            }
        mzb_replay_config rc{};
        rc.capacity = max_size; rc.min_sample_size = config.min_sample_size; rc.td_steps = config.td_steps; rc.num_unroll_steps = config.num_unroll_steps;
        rc.per_alpha = config.per_alpha; rc.per_beta = config.per_beta; rc.per_epsilon = config.per_epsilon; rc.per_beta_increment = config.per_beta_increment;
        rc.action_planes_initial = planes.empty() ? nullptr : planes.data();
        K1_ = config.num_unroll_steps + 1;
        alpha_ = config.per_alpha; epsilon_ = config.per_epsilon;
        std::lock_guard<std::mutex> lk(engine_->call_mutex());
        mzb200::check(mzb_replay_create(engine_->handle(), &rc, &r_), "mzb_replay_create");
        flat_ = engine_->shapes().initial_flat_size;
    }
    PrioritizedReplayBuffer(const muzero_config::MuZeroConfig& config, int max_size, const std::string& name, VPRNetModel& model)
        : PrioritizedReplayBuffer(config, max_size, name, model.engine()) {}
    ~PrioritizedReplayBuffer() { mzb_replay_destroy(r_); }
    PrioritizedReplayBuffer(const PrioritizedReplayBuffer&) = delete;
    PrioritizedReplayBuffer& operator=(const PrioritizedReplayBuffer&) = delete;

    bool can_sample() const { std::lock_guard<std::mutex> lk(m_); return mzb_replay_can_sample(r_) != 0; }
    int size() const { std::lock_guard<std::mutex> lk(m_); return mzb_replay_size(r_); }

    void save_game_history(const types::GameHistory& h) {
        const int L = (int)h.root_values.size();
        if (L < 1 || (int)h.observation_history.size() != L + 1 || (int)h.action_history.size() != L + 1 || (int)h.reward_history.size() != L + 1 ||
            (int)h.to_play_history.size() != L + 1 || (int)h.child_visits.size() != L)
            detail::fatal("PrioritizedReplayBuffer::save_game_history: a history of L searched positions holds L + 1 observations, actions, rewards and players");
        std::vector<float> obs;
        for (const auto& o : h.observation_history) obs.insert(obs.end(), o.begin(), o.end());
        std::vector<double> cv;
        for (const auto& v : h.child_visits) {
            if ((int)v.size() != A_) detail::fatal("PrioritizedReplayBuffer::save_game_history: child_visits row is not one entry per action");
            cv.insert(cv.end(), v.begin(), v.end());
        }
        mzb_game_history g{L, obs.data(), h.action_history.data(), h.reward_history.data(), h.to_play_history.data(), h.root_values.data(), cv.data()};
        std::lock_guard<std::mutex> lk(m_);
        {
            std::lock_guard<std::mutex> el(engine_->call_mutex());
            mzb200::check(mzb_replay_save_game_history(r_, &g), "mzb_replay_save_game_history");
        }
        host_[mzb_replay_num_histories(r_)] = h;
        prune();
    }

    template <class Rng>
    types::Batch sample(Rng& rng, int batch_size) {
        types::Batch b;
        b.priorities.resize(batch_size); b.indices.resize(batch_size); b.gradient_scale.resize(batch_size);
        b.actions.resize((size_t)batch_size * K1_); b.target_rewards.resize((size_t)batch_size * K1_); b.target_values.resize((size_t)batch_size * K1_);
        b.target_policies.resize((size_t)batch_size * K1_ * A_); b.stacked_observations.resize((size_t)batch_size * flat_);
        mzb_batch_out o{0, b.priorities.data(), b.indices.data(), b.actions.data(), b.stacked_observations.data(), b.target_rewards.data(),
                        b.target_values.data(), b.target_policies.data(), b.gradient_scale.data()};
        std::lock_guard<std::mutex> lk(m_);
        std::lock_guard<std::mutex> el(engine_->call_mutex());
        mzb200::check(mzb_replay_sample(r_, &draw<Rng>, &rng, batch_size, &o), "mzb_replay_sample");
        b.num_samples = o.num_samples;
        return b;
    }

    void update_history_priorities(const std::vector<int>& indices, std::vector<double>& errors) {
        if (indices.size() != errors.size()) detail::fatal("PrioritizedReplayBuffer::update_history_priorities: one error per index");
        std::lock_guard<std::mutex> lk(m_);
        mzb200::check(mzb_replay_update_priorities(r_, indices.data(), errors.data(), (int)indices.size()), "mzb_replay_update_priorities");
        for (double& e : errors) e = std::pow(std::abs(e) + epsilon_, alpha_);   // the reference converts `errors` in place (replay_buffer.cpp:45-49,73)
    }

    template <class Rng>
    std::tuple<int, types::GameHistory> sample_game(Rng& rng) {
        std::lock_guard<std::mutex> lk(m_);
        int32_t id = 0, len = 0;
        mzb200::check(mzb_replay_sample_game(r_, &draw<Rng>, &rng, &id, &len), "mzb_replay_sample_game");
        auto it = host_.find(id);
        return {id, (len > 0 && it != host_.end()) ? it->second : types::GameHistory()};
    }

    void update_game_history(int history_id, const types::GameHistory& h) {
        std::lock_guard<std::mutex> lk(m_);
        auto it = host_.find(history_id);
        if (it == host_.end()) return;   // lost meanwhile: nothing to do (replay_buffer.h:55-60)
        std::vector<double> cv;
        for (const auto& v : h.child_visits) cv.insert(cv.end(), v.begin(), v.end());
        std::lock_guard<std::mutex> el(engine_->call_mutex());
        mzb200::check(mzb_replay_update_game_history(r_, history_id, h.root_values.data(), cv.data(), (int)h.root_values.size()), "mzb_replay_update_game_history");
        it->second = h;
    }

    mzb_replay* handle() const { return r_; }

private:
    template <class Rng>
    static uint32_t draw(void* user) { return (uint32_t)(*static_cast<Rng*>(user))(); }
    void prune() {
        const int oldest = mzb_replay_oldest_history(r_);
        for (auto it = host_.begin(); it != host_.end();) it = it->first < oldest ? host_.erase(it) : std::next(it);
    }
    std::shared_ptr<mzb200::Engine> engine_;
    std::string name_;
    int A_, K1_ = 1, flat_ = 0;
    double alpha_ = 1, epsilon_ = 0.01;
    mzb_replay* r_ = nullptr;
    std::map<int, types::GameHistory> host_;
    mutable std::mutex m_;
};
}  // namespace buffer

class SelfPlayCohort {
public:
    // games[i] plays on tree slot i of the model's engine with search seed search_seeds[i] and actor seed actor_seeds[i]
    // (the reference gives both `config.seed + actor_num`, self_play.cpp:232-233).
    SelfPlayCohort(const muzero_config::MuZeroConfig& config, VPRNetModel& model, std::vector<AbstractGame*> games,
                   const std::vector<int>& search_seeds, const std::vector<int>& actor_seeds)
        : SelfPlayCohort(config, model.engine(), std::move(games), search_seeds, actor_seeds) {
        // the cohort plays on tree slots 0 .. n-1 and re-seeds them: MCTS objects attached to the same model hold slots from 0 up too
        if (model.cohort()->members() > 0) detail::fatal("SelfPlayCohort: the model's tree slots are in use by MCTS objects; give the cohort a model of its own");
    }
    // the same over an engine the caller built itself (e.g. one with a synthetic evaluator, to compare drivers decision for decision)
    SelfPlayCohort(const muzero_config::MuZeroConfig& config, std::shared_ptr<mzb200::Engine> engine, std::vector<AbstractGame*> games,
                   const std::vector<int>& search_seeds, const std::vector<int>& actor_seeds)
        : config_(config), engine_(std::move(engine)), games_(std::move(games)), A_((int)config.action_space.size()) {
        const int n = (int)games_.size();
        if (n < 1 || n > engine_->config().max_trees) detail::fatal("SelfPlayCohort: more games than tree slots (raise num_actors)");
        // play_game cuts a long game into slices of max_history_len moves for the trajectory queue (self_play.cpp:113-116,166-176);
        // play() returns whole games, which is the same thing only without a limit
        if (config.max_history_len > 0) detail::fatal("SelfPlayCohort: max_history_len > 0 (history slices) is not supported; play() returns whole games");
        if ((int)search_seeds.size() != n || (int)actor_seeds.size() != n) detail::fatal("SelfPlayCohort: one seed pair per game");
        std::vector<float> planes;   // action_representation_initial (types.h:43) pre-evaluated at observation resolution
        if (config.action_representation_initial)
            for (types::Action a : config.action_space) {
                const types::Observation pl = config.action_representation_initial(a);
                planes.insert(planes.end(), pl.begin(), pl.end());
            }
        std::vector<int32_t> slots(n), ss(search_seeds.begin(), search_seeds.end()), as(actor_seeds.begin(), actor_seeds.end());
        for (int i = 0; i < n; ++i) slots[i] = i;
        std::lock_guard<std::mutex> call(engine_->call_mutex());
        mzb200::check(mzb_obs_ring_create(engine_->handle(), planes.empty() ? nullptr : planes.data()), "mzb_obs_ring_create");
        mzb200::check(mzb_search_seed(engine_->handle(), 0, n, ss.data()), "mzb_search_seed");
        mzb200::check(mzb_actor_seed(engine_->handle(), slots.data(), as.data(), n), "mzb_actor_seed");
    }

    // One game per slot from reset() to the end; training_step picks the temperature (self_play.cpp:104-105).
    // is_evaluator = false: an actor (opponent = Self, exploration noise on).  is_evaluator = true: temperature 0, no noise, and in
    // games of two players config.opponent_type moves for the side that is not config.muzero_player (self_play.cpp:41-56,128-143).
    std::vector<types::GameHistory> play(int training_step, bool is_evaluator = false) {
        const int n = (int)games_.size();
        const double temperature = (is_evaluator || !config_.visit_softmax_temperature) ? 0.0 : config_.visit_softmax_temperature(training_step);
        std::vector<types::GameHistory> hist(n);
        std::vector<int32_t> live(n), zeros(n, 0);
        const size_t frame = (size_t)config_.observation_shape.c * config_.observation_shape.h * config_.observation_shape.w;
        std::vector<float> frames((size_t)n * frame);
        for (int i = 0; i < n; ++i) {   // self_play.cpp:88-95
            live[i] = i;
            const types::Observation o = games_[i]->reset();
            if (o.size() != frame) detail::fatal("AbstractGame::reset: observation size differs from observation_shape");
            std::copy(o.begin(), o.end(), frames.begin() + (size_t)i * frame);
            hist[i].action_history.push_back(0);
            hist[i].observation_history.push_back(o);
            hist[i].reward_history.push_back(0);
            hist[i].to_play_history.push_back(games_[i]->to_play());
            hist[i].legal_actions.push_back(games_[i]->legal_actions());
        }
        {
            std::lock_guard<std::mutex> call(engine_->call_mutex());
            mzb200::check(mzb_obs_ring_reset(engine_->handle(), live.data(), n), "mzb_obs_ring_reset");
            mzb200::check(mzb_obs_ring_push(engine_->handle(), live.data(), frames.data(), zeros.data(), n), "mzb_obs_ring_push");
        }
        // We always play against ourself if we are an actor or in 1 player games (self_play.cpp:108-109)
        const types::OpponentTypes opponent = (!is_evaluator || config_.num_players == 1) ? types::OpponentTypes::Self : config_.opponent_type;
        while (!live.empty()) {
            const int n_live = (int)live.size();
            // who moves by search and who is moved by the evaluation opponent (self_play.cpp:128-143)
            std::vector<int32_t> srch, opp, row_of(n, -1);
            for (int g : live) {
                if (opponent == types::OpponentTypes::Self || games_[g]->to_play() == config_.muzero_player) { row_of[g] = (int)srch.size(); srch.push_back(g); }
                else opp.push_back(g);
            }
            const int m = (int)srch.size();
            std::vector<uint32_t> mask(m, 0);
            std::vector<int32_t> to_play(m), n_children(m), depth(m), child_actions((size_t)m * A_), actions(m);
            std::vector<uint8_t> noise(m, is_evaluator ? 0 : 1);
            std::vector<double> root(m), rel((size_t)m * A_), pol((size_t)m * A_), temps(m, temperature);
            for (int k = 0; k < m; ++k) {
                mask[k] = mzb200::legal_mask_of(hist[srch[k]].legal_actions.back(), A_);
                to_play[k] = games_[srch[k]]->to_play();
            }
            mzb_search_result out{};
            out.root_value = root.data(); out.max_tree_depth = depth.data(); out.n_children = n_children.data();
            out.child_actions = child_actions.data(); out.children_relative_visit = rel.data(); out.emperical_policy = pol.data();
            if (m > 0) {   // get_stacked_observations(-1, ...) + mcts.run + select_action for every searching game (self_play.cpp:119-133)
                std::lock_guard<std::mutex> call(engine_->call_mutex());
                mzb200::check(mzb_search_batch_ring(engine_->handle(), srch.data(), mask.data(), to_play.data(), noise.data(), m, &out),
                              "mzb_search_batch_ring");
                mzb200::check(mzb_select_actions(engine_->handle(), srch.data(), &out, m, temps.data(), actions.data()), "mzb_select_actions");
                ++moves_;
            }
            // select_opponent_action (self_play.cpp:41-56)
            std::vector<types::Action> opp_action(n, types::InvalidAction);
            if (!opp.empty() && opponent == types::OpponentTypes::Random) {
                std::vector<int32_t> hi(opp.size()), pick(opp.size());
                for (size_t k = 0; k < opp.size(); ++k) hi[k] = (int32_t)hist[opp[k]].legal_actions.back().size() - 1;
                std::lock_guard<std::mutex> call(engine_->call_mutex());
                mzb200::check(mzb_actor_uniform_int(engine_->handle(), opp.data(), hi.data(), (int)opp.size(), pick.data()), "mzb_actor_uniform_int");
                for (size_t k = 0; k < opp.size(); ++k) opp_action[opp[k]] = hist[opp[k]].legal_actions.back()[pick[k]];
            } else {
                for (int g : opp) opp_action[g] = opponent == types::OpponentTypes::Expert ? games_[g]->expert_action() : games_[g]->human_to_action();
            }
            std::vector<int32_t> next_live, pushed_actions;
            std::vector<float> next_frames;
            for (int li = 0; li < n_live; ++li) {
                const int g = live[li], k = row_of[g];
                types::GameHistory& h = hist[g];
                // only searched moves store root statistics (self_play.cpp:131-133)
                if (k >= 0) h.store_search_statistics(root[k], std::vector<double>(pol.begin() + (size_t)k * A_, pol.begin() + (size_t)(k + 1) * A_));
                const types::Action played = k >= 0 ? actions[k] : opp_action[g];
                const types::StepReturn st = games_[g]->step(played);
                // the reference's own termination rule, kept verbatim (self_play.cpp:149-150): with max_moves > 0 a game also ends
                // while FEWER than max_moves searches have been stored
                const bool done = st.done || (config_.max_moves > 0 && (int)h.root_values.size() < config_.max_moves);
                h.action_history.push_back(played);
                h.observation_history.push_back(st.observation);
                h.reward_history.push_back(st.reward);
                h.to_play_history.push_back(games_[g]->to_play());
                h.legal_actions.push_back(games_[g]->legal_actions());
                if (done) continue;
                if (st.observation.size() != frame) detail::fatal("AbstractGame::step: observation size differs from observation_shape");
                next_live.push_back(g);
                pushed_actions.push_back(played);
                next_frames.insert(next_frames.end(), st.observation.begin(), st.observation.end());
            }
            if (!next_live.empty()) {
                std::lock_guard<std::mutex> call(engine_->call_mutex());
                mzb200::check(mzb_obs_ring_push(engine_->handle(), next_live.data(), next_frames.data(), pushed_actions.data(), (int)next_live.size()),
                              "mzb_obs_ring_push");
            }
            live.swap(next_live);
        }
        return hist;
    }
    long cohort_moves() const { return moves_; }   // lock-step searches run so far

private:
    muzero_config::MuZeroConfig config_;
    std::shared_ptr<mzb200::Engine> engine_;
    std::vector<AbstractGame*> games_;
    int A_;
    long moves_ = 0;
};

// reanalyze (self_play.cpp:186-225) for a cohort of sampled histories: the reference re-searches the positions of ONE history
// one after another on one actor's MCTS object; here history r lives on tree slot r (so its searches consume that slot's RNG
// stream in the reference's order) and position i of EVERY history is searched in one lock-step batch.  Frames are replayed
// through the observation ring, which rebuilds get_stacked_observations(i, ...) on the device.
class ReanalyzeCohort {
public:
    ReanalyzeCohort(const muzero_config::MuZeroConfig& config, VPRNetModel& model, const std::vector<int>& search_seeds)
        : ReanalyzeCohort(config, model.engine(), search_seeds) {
        if (model.cohort()->members() > 0) detail::fatal("ReanalyzeCohort: the model's tree slots are in use by MCTS objects; give the cohort a model of its own");
    }
    ReanalyzeCohort(const muzero_config::MuZeroConfig& config, std::shared_ptr<mzb200::Engine> engine, const std::vector<int>& search_seeds)
        : config_(config), engine_(std::move(engine)), A_((int)config.action_space.size()), n_((int)search_seeds.size()) {
        if (n_ < 1 || n_ > engine_->config().max_trees) detail::fatal("ReanalyzeCohort: more histories than tree slots");
        std::vector<float> planes;
        if (config.action_representation_initial)
            for (types::Action a : config.action_space) {
                const types::Observation pl = config.action_representation_initial(a);
                planes.insert(planes.end(), pl.begin(), pl.end());
            }
        std::vector<int32_t> ss(search_seeds.begin(), search_seeds.end());
        std::lock_guard<std::mutex> call(engine_->call_mutex());
        mzb200::check(mzb_obs_ring_create(engine_->handle(), planes.empty() ? nullptr : planes.data()), "mzb_obs_ring_create");
        mzb200::check(mzb_search_seed(engine_->handle(), 0, n_, ss.data()), "mzb_search_seed");
    }

    // histories[r] is updated in place: root_values[i] and child_visits[i] of every stored position (self_play.cpp:213-214)
    void reanalyze(std::vector<types::GameHistory*>& histories) {
        if ((int)histories.size() > n_) detail::fatal("ReanalyzeCohort: more histories than seeded slots");
        const size_t frame = (size_t)config_.observation_shape.c * config_.observation_shape.h * config_.observation_shape.w;
        std::vector<int32_t> all(histories.size());
        size_t longest = 0;
        for (size_t r = 0; r < histories.size(); ++r) { all[r] = (int32_t)r; longest = std::max(longest, histories[r]->root_values.size()); }
        {
            std::lock_guard<std::mutex> call(engine_->call_mutex());
            mzb200::check(mzb_obs_ring_reset(engine_->handle(), all.data(), (int)all.size()), "mzb_obs_ring_reset");
        }
        for (size_t i = 0; i < longest; ++i) {
            std::vector<int32_t> live, led_by;
            std::vector<float> frames;
            std::vector<uint32_t> mask;
            std::vector<int32_t> to_play;
            for (size_t r = 0; r < histories.size(); ++r) {
                const types::GameHistory& h = *histories[r];
                if (i >= h.root_values.size()) continue;
                if (h.observation_history[i].size() != frame) detail::fatal("ReanalyzeCohort: observation size differs from observation_shape");
                live.push_back((int32_t)r);
                led_by.push_back(h.action_history[i]);
                frames.insert(frames.end(), h.observation_history[i].begin(), h.observation_history[i].end());
                mask.push_back(mzb200::legal_mask_of(h.legal_actions[i], A_));
                to_play.push_back(h.to_play_history[i]);
            }
            const int m = (int)live.size();
            std::vector<uint8_t> noise(m, 1);
            std::vector<double> root(m), pol((size_t)m * A_);
            mzb_search_result out{};
            out.root_value = root.data(); out.emperical_policy = pol.data();
            {
                std::lock_guard<std::mutex> call(engine_->call_mutex());
                mzb200::check(mzb_obs_ring_push(engine_->handle(), live.data(), frames.data(), led_by.data(), m), "mzb_obs_ring_push");
                mzb200::check(mzb_search_batch_ring(engine_->handle(), live.data(), mask.data(), to_play.data(), noise.data(), m, &out),
                              "mzb_search_batch_ring");
            }
            for (int k = 0; k < m; ++k) {
                types::GameHistory& h = *histories[live[k]];
                h.root_values[i] = root[k];
                h.child_visits[i].assign(pol.begin() + (size_t)k * A_, pol.begin() + (size_t)(k + 1) * A_);
            }
        }
    }

private:
    muzero_config::MuZeroConfig config_;
    std::shared_ptr<mzb200::Engine> engine_;
    int A_, n_;
};

}  // namespace muzero_cpp
