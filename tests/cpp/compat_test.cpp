// compat_test.cpp — the reference's own names (muzero_cpp::algorithm::MCTS, Evaluator, VPRNetEvaluator, DeviceManager,
// VPRNetModel, MuZeroConfig) through include/muzero_cpp_mzb200.hpp -> libmzb200.so, used the way the reference uses them:
//   (1) tests/mcts_test.cpp's TestEvaluator + MCTS(config, seed, evaluator).run(...) against the values the UNMODIFIED reference
//       returns for the same calls (oracle/_ref reproduces them, tests/test_oracle_cpu.py);
//   (2) VPRNetModel batch inference against the reference's VPRNetModel outputs (tests/golden/net_tiny.mzbt);
//   (3) actor threads blocking on VPRNetEvaluator one-request calls (vprnet_evaluator.cpp's contract);
//   (4) actor threads calling MCTS::run over a VPRNetEvaluator, against one direct batched engine search.
// Usage: compat_test <tests/golden dir>
#define MZB200_FACADE_THROW
#include <atomic>
#include <cmath>
#include <cstdio>
#include <thread>

#include "muzero_cpp_mzb200.hpp"
#include "../../oracle/mzbt_io.h"   // test infrastructure: the golden container

using namespace muzero_cpp;
using namespace muzero_cpp::types;

static int failures = 0;
#define EXPECT(cond, ...) do { if (!(cond)) { ++failures; std::printf("FAIL %s:%d: ", __FILE__, __LINE__); std::printf(__VA_ARGS__); std::printf("\n"); } } while (0)

// tests/mcts_test.cpp:16-27
class TestEvaluator : public Evaluator {
public:
    VPRNetModel::InferenceOutputs InitialInference(const Observation&) override { return {1, 1, {0.01, 0.2, 0.3, 0.1, 0.5, 0.25}, {0}}; }
    VPRNetModel::InferenceOutputs RecurrentInference(const Observation&, Action action) override {
        return {1, (double)action, {0.01, 0.2, 0.3, 0.1, 0.5, 0.25}, {0}};
    }
};

struct Golden { double root_value; int depth; std::vector<Action> actions; std::vector<int> visits; std::vector<double> policy; };

static void expect_return(const char* name, const MCTSReturn& r, const Golden& g) {
    EXPECT(r.root_value == g.root_value, "%s: root_value %.17g != %.17g", name, r.root_value, g.root_value);
    EXPECT(r.max_tree_depth == g.depth, "%s: depth %d != %d", name, r.max_tree_depth, g.depth);
    EXPECT(r.child_actions == g.actions, "%s: child order differs", name);
    int total = 0;
    for (int v : g.visits) total += v;
    for (size_t i = 0; i < g.visits.size() && i < r.children_relative_visit.size(); ++i)
        EXPECT(r.children_relative_visit[i] == (double)g.visits[i] / total, "%s: relative visit %zu", name, i);
    EXPECT(r.emperical_policy.size() == g.policy.size(), "%s: policy size", name);
    for (size_t i = 0; i < g.policy.size() && i < r.emperical_policy.size(); ++i)
        EXPECT(std::fabs(r.emperical_policy[i] - g.policy[i]) < 1e-12, "%s: policy[%zu] %.17g != %.17g", name, i, r.emperical_policy[i], g.policy[i]);
}

static void appendix_b() {
    // tests/mcts_test.cpp:229-241
    muzero_config::MuZeroConfig config;
    config.seed = 0; config.num_players = 2; config.num_simulations = 10;
    config.dirichlet_alpha = 0.3; config.dirichlet_epsilon = 0.25; config.pb_c_base = 1; config.pb_c_init = 0; config.discount = 0.5;
    config.value_lowerbound = 0; config.value_upperbound = 1;
    config.action_space = {0, 1, 2, 3, 4, 5};
    const Observation obs{0};
    {
        algorithm::MCTS mcts(config, 0, std::make_shared<TestEvaluator>());
        expect_return("G1 run 1", mcts.run(obs, {0, 1, 4, 5}, 0, false), {2.5531250000000001, 5, {1, 5, 4, 0}, {1, 3, 5, 1}, {0.1, 0.1, 0, 0, 0.5, 0.3}});
        expect_return("G1 run 2", mcts.run(obs, {0, 1, 4, 5}, 1, false), {2.9281250000000001, 5, {4, 1, 0, 5}, {6, 1, 0, 3}, {0, 0.1, 0, 0, 0.6, 0.3}});
    }
    {
        algorithm::MCTS mcts(config, 0, std::make_shared<TestEvaluator>());
        expect_return("G2 run 1", mcts.run(obs, {0, 1, 4, 5}, 0, true), {2.859375, 5, {1, 0, 4, 5}, {0, 1, 5, 4}, {0.1, 0, 0, 0, 0.5, 0.4}});
        expect_return("G2 run 2", mcts.run(obs, {0, 1, 4, 5}, 1, true), {2.8046875, 6, {1, 4, 0, 5}, {1, 7, 0, 2}, {0, 0.1, 0, 0, 0.7, 0.2}});
    }
    muzero_config::MuZeroConfig defaults;
    defaults.num_players = 2; defaults.value_lowerbound = -10; defaults.value_upperbound = 10;
    defaults.action_space = {0, 1, 2, 3, 4, 5};
    {
        algorithm::MCTS mcts(defaults, 7, std::make_shared<TestEvaluator>());
        expect_return("G3", mcts.run(obs, {0, 1, 2, 3, 4, 5}, 0, true),
                      {1.8830028913384134, 8, {3, 5, 2, 4, 1, 0}, {0, 13, 7, 26, 4, 0}, {0, 0.08, 0.14, 0, 0.52, 0.26}});
    }
    muzero_config::MuZeroConfig single;
    single.num_players = 1;
    single.action_space = {0, 1, 2, 3, 4, 5};
    {
        algorithm::MCTS mcts(single, 7, std::make_shared<TestEvaluator>());
        expect_return("G4", mcts.run(obs, {0, 1, 2, 3, 4, 5}, 0, true),
                      {16.064277835397309, 9, {1, 2, 5, 4, 0, 3}, {1, 2, 22, 25, 0, 0}, {0, 0.02, 0.04, 0, 0.5, 0.44}});
    }
}

// the "tiny" network of tests/golden/make_golden.py
static muzero_config::MuZeroConfig tiny_config(int actors) {
    muzero_config::MuZeroConfig c;
    c.observation_shape = {2, 3, 4};
    c.action_space = {0, 1, 2, 3, 4};
    c.action_channels = 1; c.num_players = 2; c.stacked_observations = 1;
    c.value_upperbound = 5; c.value_lowerbound = -5; c.min_reward = -3; c.max_reward = 3; c.min_value = -5; c.max_value = 5;
    c.num_simulations = 12;
    c.num_actors = actors; c.num_reanalyze_actors = 0; c.num_evaluators = 0;
    c.initial_inference_batch_size = 4; c.recurrent_inference_batch_size = 4;
    auto& n = c.network_config;
    n.resnet_channels = 16; n.representation_blocks = 1; n.dynamics_blocks = 1; n.prediction_blocks = 2;
    n.reward_reduced_channels = 4; n.policy_reduced_channels = 2; n.value_reduced_channels = 3;
    n.reward_head_layers = {8}; n.policy_head_layers = {8, 6}; n.value_head_layers = {};
    return c;
}

static const mzbt::Tensor& find(const std::vector<mzbt::Tensor>& all, const char* name) {
    for (auto& t : all) if (t.name == name) return t;
    std::printf("golden tensor %s missing\n", name);
    std::exit(2);
}
static std::vector<mzb_tensor_desc> weight_descs(const std::vector<mzbt::Tensor>& all) {
    std::vector<mzb_tensor_desc> w;
    for (auto& t : all) {
        if (t.name.rfind("w.", 0) != 0) continue;
        mzb_tensor_desc d{};
        d.name = t.name.c_str() + 2; d.data = t.as<float>(); d.ndim = (int)t.dims.size();
        for (size_t k = 0; k < t.dims.size(); ++k) d.dims[k] = t.dims[k];
        w.push_back(d);
    }
    return w;
}

static double max_abs_diff(const VPRNetModel::InferenceOutputs& o, const double* value, const double* reward, const double* policy, int A,
                           const float* hidden, int S) {
    double m = std::max(std::fabs(o.value - *value), std::fabs(o.reward - *reward));
    for (int a = 0; a < A; ++a) m = std::max(m, std::fabs(o.policy[a] - policy[a]));
    for (int k = 0; k < S; ++k) m = std::max(m, (double)std::fabs(o.encoded_state[k] - hidden[k]));
    return m;
}

static void network_paths(const std::string& golden_dir) {
    std::vector<mzbt::Tensor> g;
    if (!mzbt::load(golden_dir + "/net_tiny.mzbt", g)) { std::printf("cannot read net_tiny.mzbt\n"); std::exit(2); }
    const std::vector<mzb_tensor_desc> weights = weight_descs(g);
    const int B = 3, A = 5, S = 192, F = 60, actors = 6;
    const mzbt::Tensor &obs = find(g, "in.obs"), &a1 = find(g, "in.a1");

    // (2) VPRNetModel, fp32 check mode, against the reference's outputs (1e-4, north star)
    DeviceManager dm;
    {
        VPRNetModel model(tiny_config(actors), "cuda:0", MZB_FP32_CHECK);
        model.LoadWeights(weights);
        EXPECT(model.Device() == "cuda:0", "Device()");
        dm.AddDevice(std::move(model));
    }
    EXPECT(dm.Count() == 1, "Count()");
    std::vector<VPRNetModel::InitialInferenceInputs> in0(B);
    for (int b = 0; b < B; ++b) in0[b].stacked_observation.assign(obs.as<float>() + b * F, obs.as<float>() + (b + 1) * F);
    std::vector<VPRNetModel::InferenceOutputs> init = dm.Get(B)->InitialInference(in0);
    for (int b = 0; b < B; ++b) {
        const double d = max_abs_diff(init[b], find(g, "init.value").as<double>() + b, find(g, "init.reward").as<double>() + b,
                                      find(g, "init.policy").as<double>() + b * A, A, find(g, "init.hidden").as<float>() + b * S, S);
        EXPECT(d <= 1e-4, "InitialInference row %d differs from the reference by %g", b, d);
    }
    std::vector<VPRNetModel::RecurrentInferenceInputs> in1(B);
    for (int b = 0; b < B; ++b) { in1[b].action = a1.as<int32_t>()[b]; in1[b].encoded_state = init[b].encoded_state; }
    std::vector<VPRNetModel::InferenceOutputs> rec = dm.Get(B, 0)->RecurrentInference(in1);
    for (int b = 0; b < B; ++b) {
        const double d = max_abs_diff(rec[b], find(g, "rec1.value").as<double>() + b, find(g, "rec1.reward").as<double>() + b,
                                      find(g, "rec1.policy").as<double>() + b * A, A, find(g, "rec1.hidden").as<float>() + b * S, S);
        EXPECT(d <= 2e-4, "RecurrentInference row %d differs from the reference by %g", b, d);
    }

    // (3) VPRNetEvaluator: blocking one-request calls from many threads are answered in batches and equal the batch call
    auto evaluator = std::make_shared<VPRNetEvaluator>(&dm, 4, 1, 4, 2);
    {
        const int threads = 9;
        std::vector<std::thread> pool;
        std::atomic<int> bad{0};
        for (int i = 0; i < threads; ++i)
            pool.emplace_back([&, i] {
                const int b = i % B;
                for (int rep = 0; rep < 5; ++rep) {
                    VPRNetModel::InferenceOutputs o0 = evaluator->InitialInference(in0[b].stacked_observation);
                    VPRNetModel::InferenceOutputs o1 = evaluator->RecurrentInference(o0.encoded_state, in1[b].action);
                    if (o0.value != init[b].value || o0.policy != init[b].policy || o0.encoded_state != init[b].encoded_state) ++bad;
                    if (o1.value != rec[b].value || o1.reward != rec[b].reward || o1.policy != rec[b].policy || o1.encoded_state != rec[b].encoded_state) ++bad;
                }
            });
        for (auto& t : pool) t.join();
        EXPECT(bad == 0, "%d evaluator answers differ from the batch call", (int)bad);
        EXPECT(evaluator->requests_served() == threads * 5 * 2, "requests served %ld", evaluator->requests_served());
        EXPECT(evaluator->batches_served() < evaluator->requests_served(), "no batching happened (%ld batches)", evaluator->batches_served());
    }

    // (4) self_play.cpp:228-238: one MCTS object per actor thread over the shared evaluator, against ONE direct batched search
    //     on a second engine with the same weights and seeds
    muzero_config::MuZeroConfig config = tiny_config(actors);
    const int moves = 2;
    std::vector<std::vector<MCTSReturn>> want(moves, std::vector<MCTSReturn>(actors));
    {
        std::vector<float> planes;
        mzb200::Engine direct(detail::to_pod(config, actors, MZB_FP32_CHECK, MZB_EVAL_NETWORK, &planes), 0);
        direct.LoadWeights(weights);
        std::vector<int32_t> seeds(actors);
        for (int t = 0; t < actors; ++t) seeds[t] = config.seed + t;
        mzb200::check(mzb_search_seed(direct.handle(), 0, actors, seeds.data()), "seed");
        for (int m = 0; m < moves; ++m) {
            std::vector<float> rows((size_t)actors * F);
            std::vector<uint32_t> mask(actors);
            std::vector<int32_t> tp(actors), nch(actors), depth(actors), act(actors * A);
            std::vector<uint8_t> noise(actors, 1);
            std::vector<double> root(actors), rel(actors * A), pol(actors * A);
            for (int t = 0; t < actors; ++t) {
                for (int k = 0; k < F; ++k) rows[(size_t)t * F + k] = obs.as<float>()[((t + m) % B) * F + k];
                mask[t] = (t % 2) ? 0x1f : 0x1b; tp[t] = (t + m) % 2;
            }
            mzb_search_result out{};
            out.root_value = root.data(); out.max_tree_depth = depth.data(); out.n_children = nch.data();
            out.child_actions = act.data(); out.children_relative_visit = rel.data(); out.emperical_policy = pol.data();
            mzb200::check(mzb_search_batch(direct.handle(), rows.data(), mask.data(), tp.data(), noise.data(), actors, &out), "batch");
            for (int t = 0; t < actors; ++t) {
                want[m][t].root_value = root[t]; want[m][t].max_tree_depth = depth[t];
                want[m][t].child_actions.assign(act.begin() + t * A, act.begin() + t * A + nch[t]);
                want[m][t].children_relative_visit.assign(rel.begin() + t * A, rel.begin() + t * A + nch[t]);
                want[m][t].emperical_policy.assign(pol.begin() + t * A, pol.begin() + (t + 1) * A);
            }
        }
    }
    {
        std::vector<std::unique_ptr<algorithm::MCTS>> mcts(actors);
        for (int t = 0; t < actors; ++t) mcts[t].reset(new algorithm::MCTS(config, config.seed + t, evaluator));   // self_play.cpp:232
        std::atomic<int> bad{0};
        std::vector<std::thread> pool;
        for (int t = 0; t < actors; ++t)
            pool.emplace_back([&, t] {
                for (int m = 0; m < moves; ++m) {
                    Observation o(obs.as<float>() + ((t + m) % B) * F, obs.as<float>() + ((t + m) % B + 1) * F);
                    const std::vector<Action> legal = (t % 2) ? std::vector<Action>{0, 1, 2, 3, 4} : std::vector<Action>{0, 1, 3, 4};
                    MCTSReturn r = mcts[t]->run(o, legal, (t + m) % 2, true);
                    const MCTSReturn& w = want[m][t];
                    if (r.root_value != w.root_value || r.max_tree_depth != w.max_tree_depth || r.child_actions != w.child_actions ||
                        r.children_relative_visit != w.children_relative_visit || r.emperical_policy != w.emperical_policy) ++bad;
                }
            });
        for (auto& th : pool) th.join();
        EXPECT(bad == 0, "%d MCTS::run results differ from the direct batched search", (int)bad);
    }
}

int main(int argc, char** argv) {
    try {
        appendix_b();
        if (argc > 1) network_paths(argv[1]);
    } catch (const std::exception& e) {
        std::printf("compat_test: %s\n", e.what());
        return 1;
    }
    if (failures) { std::printf("compat_test: %d failure(s)\n", failures); return 1; }
    std::printf("compat_test: ok\n");
    return 0;
}
