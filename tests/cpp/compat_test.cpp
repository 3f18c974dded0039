// compat_test.cpp — the reference's own names (muzero_cpp::algorithm::MCTS, Evaluator, VPRNetEvaluator, DeviceManager,
// VPRNetModel, MuZeroConfig) through include/muzero_cpp_mzb200.hpp -> libmzb200.so, used the way the reference uses them:
//   (1) tests/mcts_test.cpp's TestEvaluator + MCTS(config, seed, evaluator).run(...) against the values the UNMODIFIED reference
//       returns for the same calls (oracle/_ref reproduces them, tests/test_oracle_cpu.py);
//   (2) VPRNetModel batch inference against the reference's VPRNetModel outputs (tests/golden/net_tiny.mzbt);
//   (3) actor threads blocking on VPRNetEvaluator one-request calls (vprnet_evaluator.cpp's contract);
//   (4) actor threads calling MCTS::run over a VPRNetEvaluator, against one direct batched engine search;
//   (5) SelfPlayCohort (play_game for a cohort of AbstractGames) against the reference's thread-per-actor loop.
//   (0) tests/mcts_test.cpp's five unit tests (backpropogate, select_child, ucb_score, MinMaxStats, MCTSNode) on the device step.
// Usage: compat_test <tests/golden dir>
#define MZB200_FACADE_THROW
#include <atomic>
#include <cmath>
#include <cstdio>
#include <random>
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

// ---------------------------------------------------------------------------------------------------------------------
// (0) tests/mcts_test.cpp:30-224, function by function, with the reference's own expected values.  The MCTSNode records are
//     assembled by hand exactly as there; select_child / ucb_score / backpropogate run on the DEVICE step (mzb_debug_*).
#define REQUIRE_NEAR(a, b, tol) EXPECT(std::fabs((a) - (b)) <= (tol), "%s = %.9g, expected %.9g", #a, (double)(a), (double)(b))
#define REQUIRE_EQUAL(a, b) EXPECT((a) == (b), "%s != %s", #a, #b)

static void mcts_backpropogate_test(const muzero_config::MuZeroConfig& config) {   // tests/mcts_test.cpp:30-97
    using algorithm::MCTSNode;
    for (int players = 2; players >= 1; --players) {
        muzero_config::MuZeroConfig c = config;
        c.num_players = players;
        algorithm::MCTS mcts_(c, c.seed, std::make_shared<TestEvaluator>());
        MCTSNode root(-1, 0);
        std::vector<Action> actions{0, 1, 4, 5};
        Player player = 0;
        std::vector<double> policy_probs{0.01, 0.2, 0.3, 0.1, 0.5, 0.4};
        Observation encoded_state;
        MCTSNode* node = &root;
        std::vector<MCTSNode*> search_path{node};
        for (int i = 0; i < 4; ++i) {
            node->expand(actions, player, 0.1, policy_probs, encoded_state);
            player = (player + 1) % players;
            node = &node->children[0];
            node->reward = 0.1;
            search_path.push_back(node);
        }
        node->expand(actions, player, 0.1, policy_probs, encoded_state);
        const std::vector<double> expected_values = players == 2 ? std::vector<double>{1, -0.4, 0.3, -0.05, 0.125}
                                                                 : std::vector<double>{1, 0.6, 0.4, 0.3, 0.25};
        mcts_.backpropogate(search_path, 1, player);
        for (int i = (int)search_path.size() - 1; i >= 0; --i) {
            REQUIRE_NEAR(search_path[i]->value(), expected_values[expected_values.size() - i - 1], 0.001);
            REQUIRE_EQUAL(search_path[i]->visit_count, 1);
        }
    }
}

static void mcts_select_child_test(const muzero_config::MuZeroConfig& config) {    // tests/mcts_test.cpp:100-135
    using algorithm::MCTSNode;
    algorithm::MCTS mcts_(config, config.seed, std::make_shared<TestEvaluator>());
    MCTSNode root(-1, 0);
    std::vector<Action> actions{0, 1, 4, 5};
    std::vector<double> policy_probs{0.01, 0.2, 0.3, 0.1, 0.5, 0.4};
    root.expand(actions, 0, 10, policy_probs, Observation());
    for (auto& ch : root.children) { ch.reward = 1; ch.visit_count = 1; ch.value_sum = 0.5; }
    // Root has no visits, so prior has no influence: every score ties and the seeded shuffle decides (mcts_test.cpp:118-119)
    MCTSNode* child = mcts_.select_child(&root);
    REQUIRE_EQUAL(child->action, 0);
    std::vector<Action> seen;
    for (auto& ch : root.children) seen.push_back(ch.action);
    std::sort(seen.begin(), seen.end());
    EXPECT(seen == actions, "select_child lost or duplicated a child while shuffling");
    root.visit_count += 1;
    child = mcts_.select_child(&root);
    REQUIRE_EQUAL(child->action, 4);
    for (auto& ch : root.children)
        if (ch.action == 1) { ch.value_sum -= 100; ch.visit_count += 100; }
    child = mcts_.select_child(&root);
    REQUIRE_EQUAL(child->action, 1);
}

static void mcts_ucb_test(const muzero_config::MuZeroConfig& config) {             // tests/mcts_test.cpp:138-153
    using algorithm::MCTSNode;
    algorithm::MCTS mcts_(config, config.seed, std::make_shared<TestEvaluator>());
    MCTSNode root(-1, 0);
    std::vector<double> policy_probs{0.01, 0.2, 0.3, 0.1, 0.5, 0.4};
    root.visit_count = 2;
    root.expand({0, 1, 4, 5}, 0, 10, policy_probs, Observation());
    root.children[0].reward = 2;
    root.children[0].value_sum = -8;
    root.children[0].visit_count = 10;
    REQUIRE_NEAR(mcts_.ucb_score(&root, &root.children[0]), 2.401782, 0.001);
    REQUIRE_EQUAL(root.children[0].action, 0);   // ucb_score neither shuffles nor draws
}

// MinMaxStats (mcts.h:79-112, tests/mcts_test.cpp:156-188) as the tree kernels hold it: normalize(v) is the value term of the
// pUCT score of a once-visited child with reward v under an unvisited parent (the prior term vanishes, mcts.cpp:130-139);
// update(v) is one backup through a lone root whose expansion reward is v and whose value is 0 (mcts.cpp:99).
struct DeviceMinMax {
    mzb200::Engine eng;
    double lo, hi, init_lo, init_hi;
    DeviceMinMax(const mzb_config& pod, double minimum, double maximum) : eng(pod, 0), lo(minimum), hi(maximum), init_lo(minimum), init_hi(maximum) {}
    void reset() { lo = init_lo; hi = init_hi; }
    double normalize(double v) {
        const int32_t meta[8] = {0, -1, 1, 0, 1, 0, -1, 0};
        const double prior[2] = {0, 0.5}, vsum[2] = {0, 0}, rew[2] = {0, v};
        mzb_debug_tree t{2, meta, prior, vsum, rew, lo, hi, 0, 1};
        mzb200::check(mzb_debug_tree_write(eng.handle(), 0, &t), "write");
        int32_t chosen;
        std::vector<double> scores(eng.config().num_actions);
        mzb200::check(mzb_debug_select_child(eng.handle(), 0, 1, &chosen, scores.data()), "select");
        return scores[0];
    }
    void update(double v) {
        const int32_t meta[4] = {0, -1, -1, 0}, path[1] = {0};
        const double zero[1] = {0};
        mzb_debug_tree t{1, meta, zero, zero, zero, lo, hi, 0, 0};
        mzb200::check(mzb_debug_tree_write(eng.handle(), 0, &t), "write");
        const std::vector<double> policy(eng.config().num_actions, 0.0);
        mzb200::check(mzb_debug_backpropagate(eng.handle(), 0, path, 1, 0.0, v, policy.data()), "backup");
        double mm[2];
        mzb200::check(mzb_debug_tree_read(eng.handle(), 0, 1, nullptr, nullptr, nullptr, nullptr, mm), "read");
        lo = mm[0]; hi = mm[1];
    }
};

static void minmax_stats_test(const muzero_config::MuZeroConfig& config) {
    muzero_config::MuZeroConfig c = config;
    c.num_players = 1; c.discount = 1.0;
    const mzb_config pod = detail::to_pod(c, 1, MZB_FP32_CHECK, MZB_EVAL_HOST, nullptr);
    {   // default bounds
        DeviceMinMax minmax_stats(pod, INF_D, NINF_D);
        REQUIRE_EQUAL(minmax_stats.normalize(10), 10);
        minmax_stats.update(2);
        REQUIRE_EQUAL(minmax_stats.normalize(10), 10);
        minmax_stats.update(12);
        REQUIRE_EQUAL(minmax_stats.normalize(10), 0.8);
        REQUIRE_EQUAL(minmax_stats.normalize(2), 0.0);
        REQUIRE_EQUAL(minmax_stats.normalize(12), 1.0);
        minmax_stats.reset();
        REQUIRE_EQUAL(minmax_stats.normalize(10), 10);
        minmax_stats.update(2);
        REQUIRE_EQUAL(minmax_stats.normalize(10), 10);
    }
    {   // bounds given
        DeviceMinMax minmax_stats(pod, 0, 1);
        REQUIRE_EQUAL(minmax_stats.normalize(0.5), 0.5);
        REQUIRE_EQUAL(minmax_stats.normalize(0), 0.0);
        REQUIRE_EQUAL(minmax_stats.normalize(1), 1.0);
        minmax_stats.update(0.5);
        REQUIRE_EQUAL(minmax_stats.normalize(0.5), 0.5);
        REQUIRE_EQUAL(minmax_stats.normalize(0), 0.0);
        REQUIRE_EQUAL(minmax_stats.normalize(1), 1.0);
    }
}

static void mctsnode_test() {                                                      // tests/mcts_test.cpp:191-213 (bookkeeping part)
    algorithm::MCTSNode root(-1, 0);
    EXPECT(!root.is_expanded(), "fresh node is expanded");
    REQUIRE_EQUAL(root.value(), 0);
    std::vector<Action> actions{0, 1, 4, 5};
    std::vector<double> policy_probs{0.01, 0.2, 0.3, 0.1, 0.5, 0.25};
    root.expand(actions, 0, 10, policy_probs, Observation());
    root.value_sum += 10;
    root.visit_count += 2;
    REQUIRE_EQUAL(root.children.size(), actions.size());
    REQUIRE_EQUAL(root.reward, 10);
    REQUIRE_EQUAL(root.value(), 5);
    EXPECT(root.is_expanded(), "expanded node is not expanded");
    for (size_t i = 0; i < root.children.size(); ++i) {
        REQUIRE_EQUAL(root.children[i].prior, policy_probs[actions[i]]);
        REQUIRE_EQUAL(root.children[i].action, actions[i]);
        EXPECT(!root.children[i].is_expanded(), "fresh child is expanded");
    }
}

static void reference_unit_tests() {
    muzero_config::MuZeroConfig config;   // tests/mcts_test.cpp:229-241
    config.seed = 0; config.num_players = 2; config.num_simulations = 10;
    config.dirichlet_alpha = 0.3; config.dirichlet_epsilon = 0.25; config.pb_c_base = 1; config.pb_c_init = 0; config.discount = 0.5;
    config.value_lowerbound = 0; config.value_upperbound = 1;
    config.action_space = {0, 1, 2, 3, 4, 5};
    mcts_backpropogate_test(config);
    mcts_select_child_test(config);
    mcts_ucb_test(config);
    minmax_stats_test(config);
    mctsnode_test();
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

// ---------------------------------------------------------------------------------------------------------------------
// (5) SelfPlayCohort against the reference's actor pattern.  A toy two-player game with ragged lengths and changing legal moves.
class DriftGame : public AbstractGame {
public:
    explicit DriftGame(int id) : id_(id) {}
    Observation reset() override { k_ = 0; return obs(); }
    StepReturn step(Action a) override {
        ++k_;
        last_ = a;
        return {obs(), 0.25 * a - 0.5, k_ >= 3 + id_ % 3};
    }
    Player to_play() const override { return (k_ + id_) % 2; }
    std::vector<Action> legal_actions() const override {
        std::vector<Action> l;
        for (Action a = 0; a < 5; ++a) if (!(k_ % 2 == 1 && a == (k_ + id_) % 5)) l.push_back(a);
        return l;
    }
    Action expert_action() override { return legal_actions()[0]; }
    Action human_to_action() override { return legal_actions()[0]; }
    std::string action_to_string(Action a) const override { return std::to_string(a); }

private:
    Observation obs() const {
        Observation o(24);
        for (int i = 0; i < 12; ++i) { const double x = (id_ * 7 + k_ * 3 + i + last_) * 0.137; o[i] = (float)(x - std::floor(x)); }
        for (int i = 12; i < 24; ++i) o[i] = to_play() ? 1.f : -1.f;
        return o;
    }
    int id_, k_ = 0, last_ = 0;
};

// GameHistory::get_stacked_observations(-1, ...) as the reference's actor computes it on the host (types.cpp:16-42), restated for the check
static Observation stacked(const GameHistory& h, const muzero_config::MuZeroConfig& c) {
    const int step = (int)h.observation_history.size() - 1, flat = c.observation_shape.h * c.observation_shape.w;
    Observation out = h.observation_history[step];
    for (int past = step - 1; past >= step - c.stacked_observations; --past) {
        if (past >= 0) {
            out.insert(out.end(), h.observation_history[past].begin(), h.observation_history[past].end());
            const Observation planes = c.action_representation_initial(h.action_history[past + 1]);
            out.insert(out.end(), planes.begin(), planes.end());
        } else {
            out.insert(out.end(), (size_t)(c.observation_shape.c + c.action_channels) * flat, 0.f);
        }
    }
    return out;
}

static void self_play_cohort(const std::string& golden_dir) {
    std::vector<mzbt::Tensor> g;
    mzbt::load(golden_dir + "/net_tiny.mzbt", g);
    const std::vector<mzb_tensor_desc> weights = weight_descs(g);
    const int n = 6;
    muzero_config::MuZeroConfig config = tiny_config(n);
    config.seed = 11;
    config.action_representation_initial = [](Action a) { return Observation(12, (float)a / 5.f); };
    config.visit_softmax_temperature = [](int step) { return step < 100 ? 1.0 : 0.25; };

    // (a) the reference's pattern: one thread, one MCTS, one std::mt19937 per game (self_play.cpp:84-183, 228-238)
    std::vector<GameHistory> want(n);
    {
        DeviceManager dm;
        VPRNetModel model(config, "cuda:0", MZB_FP32_CHECK);
        model.LoadWeights(weights);
        dm.AddDevice(std::move(model));
        auto evaluator = std::make_shared<VPRNetEvaluator>(&dm, 4, 1, 4, 1);
        std::vector<std::unique_ptr<algorithm::MCTS>> mcts(n);
        for (int i = 0; i < n; ++i) mcts[i].reset(new algorithm::MCTS(config, config.seed + i, evaluator));
        std::vector<std::thread> actors;
        for (int i = 0; i < n; ++i)
            actors.emplace_back([&, i] {
                DriftGame game(i);
                std::mt19937 rng(config.seed + i);
                GameHistory& h = want[i];
                h.action_history.push_back(0);
                h.observation_history.push_back(game.reset());
                h.reward_history.push_back(0);
                h.to_play_history.push_back(game.to_play());
                h.legal_actions.push_back(game.legal_actions());
                const double temperature = config.visit_softmax_temperature(0);
                bool done = false;
                while (!done) {
                    MCTSReturn st = mcts[i]->run(stacked(h, config), game.legal_actions(), game.to_play(), true);
                    std::vector<double> w = st.children_relative_visit;
                    for (double& v : w) v = std::pow(v, 1.0 / temperature);
                    std::discrete_distribution<> d(w.begin(), w.end());
                    const Action a = st.child_actions[d(rng)];
                    h.store_search_statistics(st.root_value, st.emperical_policy);
                    StepReturn sr = game.step(a);
                    done = sr.done;
                    h.action_history.push_back(a);
                    h.observation_history.push_back(sr.observation);
                    h.reward_history.push_back(sr.reward);
                    h.to_play_history.push_back(game.to_play());
                    h.legal_actions.push_back(game.legal_actions());
                }
                mcts[i].reset();   // a finished actor leaves the cohort
            });
        for (auto& t : actors) t.join();
    }
    // (b) the cohort driver on a second model, its weights read from the checkpoint file the reference wrote for them
    //     (VPRNetModel::LoadCheckpoint(path) appends ".pt", vprnet.cpp:71-74)
    VPRNetModel model(config, "cuda:0", MZB_FP32_CHECK);
    model.LoadCheckpoint(golden_dir + "/checkpoint_tiny");
    std::vector<std::unique_ptr<DriftGame>> owned;
    std::vector<AbstractGame*> games;
    std::vector<int> seeds;
    for (int i = 0; i < n; ++i) { owned.emplace_back(new DriftGame(i)); games.push_back(owned.back().get()); seeds.push_back(config.seed + i); }
    SelfPlayCohort cohort(config, model, games, seeds, seeds);
    std::vector<GameHistory> got = cohort.play(/*training_step*/ 0);
    EXPECT(cohort.cohort_moves() == 5, "lock-step searches: %ld (longest game has 5 moves)", cohort.cohort_moves());
    for (int i = 0; i < n; ++i) {
        EXPECT((int)got[i].root_values.size() == 3 + i % 3, "game %d: %zu moves", i, got[i].root_values.size());
        EXPECT(got[i].action_history == want[i].action_history, "game %d: actions differ from the actor path", i);
        EXPECT(got[i].root_values == want[i].root_values, "game %d: root values differ", i);
        EXPECT(got[i].child_visits == want[i].child_visits, "game %d: stored policies differ", i);
        EXPECT(got[i].observation_history == want[i].observation_history && got[i].reward_history == want[i].reward_history &&
               got[i].to_play_history == want[i].to_play_history && got[i].legal_actions == want[i].legal_actions, "game %d: history differs", i);
    }
    // (6) reanalyze (self_play.cpp:186-225): the reference's loop — one MCTS object per reanalyze actor re-searching its history
    //     position by position — against ReanalyzeCohort on the played histories
    {
        std::vector<GameHistory> ref_path = got, cohort_path = got;
        {
            DeviceManager dm;
            VPRNetModel m2(config, "cuda:0", MZB_FP32_CHECK);
            m2.LoadWeights(weights);
            dm.AddDevice(std::move(m2));
            auto evaluator = std::make_shared<VPRNetEvaluator>(&dm, 4, 1, 4, 1);
            std::vector<std::unique_ptr<algorithm::MCTS>> mcts(n);
            for (int i = 0; i < n; ++i) mcts[i].reset(new algorithm::MCTS(config, 500 + i, evaluator));
            std::vector<std::thread> actors;
            for (int r = 0; r < n; ++r)
                actors.emplace_back([&, r] {
                    GameHistory& h = ref_path[r];
                    for (int i = 0; i < (int)h.root_values.size(); ++i) {
                        GameHistory upto = h;   // get_stacked_observations(i, ...): the history as it stood at step i
                        upto.observation_history.resize(i + 1);
                        MCTSReturn st = mcts[r]->run(stacked(upto, config), h.legal_actions[i], h.to_play_history[i], true);
                        h.root_values[i] = st.root_value;
                        h.child_visits[i] = st.emperical_policy;
                    }
                    mcts[r].reset();
                });
            for (auto& t : actors) t.join();
        }
        VPRNetModel m3(config, "cuda:0", MZB_FP32_CHECK);
        m3.LoadWeights(weights);
        std::vector<int> rseeds;
        for (int i = 0; i < n; ++i) rseeds.push_back(500 + i);
        ReanalyzeCohort re(config, m3, rseeds);
        std::vector<GameHistory*> ptrs;
        for (auto& h : cohort_path) ptrs.push_back(&h);
        re.reanalyze(ptrs);
        for (int r = 0; r < n; ++r) {
            EXPECT(cohort_path[r].root_values == ref_path[r].root_values, "reanalyze history %d: root values differ", r);
            EXPECT(cohort_path[r].child_visits == ref_path[r].child_visits, "reanalyze history %d: policies differ", r);
            EXPECT(cohort_path[r].root_values != got[r].root_values, "reanalyze history %d: nothing was re-searched", r);
        }
    }
    // (8) evaluator games (self_play.cpp:104-143 with is_evaluator): temperature 0, no noise, and a Random opponent that moves
    //     for player 1 by std::uniform_int_distribution draws on the actor's own generator
    {
        muzero_config::MuZeroConfig ec = config;
        ec.opponent_type = OpponentTypes::Random;
        ec.muzero_player = 0;
        std::vector<GameHistory> ref_games(n);
        {
            DeviceManager dm;
            VPRNetModel m2(ec, "cuda:0", MZB_FP32_CHECK);
            m2.LoadWeights(weights);
            dm.AddDevice(std::move(m2));
            auto evaluator = std::make_shared<VPRNetEvaluator>(&dm, 4, 1, 4, 1);
            std::vector<std::unique_ptr<algorithm::MCTS>> mcts(n);
            for (int i = 0; i < n; ++i) mcts[i].reset(new algorithm::MCTS(ec, 700 + i, evaluator));
            std::vector<std::thread> actors;
            for (int i = 0; i < n; ++i)
                actors.emplace_back([&, i] {
                    DriftGame game(i);
                    std::mt19937 rng(700 + i);
                    GameHistory& h = ref_games[i];
                    h.action_history.push_back(0);
                    h.observation_history.push_back(game.reset());
                    h.reward_history.push_back(0);
                    h.to_play_history.push_back(game.to_play());
                    h.legal_actions.push_back(game.legal_actions());
                    bool done = false;
                    while (!done) {
                        Action a;
                        if (game.to_play() == ec.muzero_player) {
                            MCTSReturn st = mcts[i]->run(stacked(h, ec), game.legal_actions(), game.to_play(), false);
                            const auto best = std::max_element(st.children_relative_visit.begin(), st.children_relative_visit.end());
                            a = st.child_actions[best - st.children_relative_visit.begin()];
                            h.store_search_statistics(st.root_value, st.emperical_policy);
                        } else {
                            const std::vector<Action> legal = game.legal_actions();
                            std::uniform_int_distribution<> d(0, (int)legal.size() - 1);
                            a = legal[d(rng)];
                        }
                        StepReturn sr = game.step(a);
                        done = sr.done;
                        h.action_history.push_back(a);
                        h.observation_history.push_back(sr.observation);
                        h.reward_history.push_back(sr.reward);
                        h.to_play_history.push_back(game.to_play());
                        h.legal_actions.push_back(game.legal_actions());
                    }
                    mcts[i].reset();
                });
            for (auto& t : actors) t.join();
        }
        VPRNetModel m3(ec, "cuda:0", MZB_FP32_CHECK);
        m3.LoadWeights(weights);
        std::vector<std::unique_ptr<DriftGame>> eg;
        std::vector<AbstractGame*> eptr;
        std::vector<int> eseeds;
        for (int i = 0; i < n; ++i) { eg.emplace_back(new DriftGame(i)); eptr.push_back(eg.back().get()); eseeds.push_back(700 + i); }
        SelfPlayCohort ev(ec, m3, eptr, eseeds, eseeds);
        std::vector<GameHistory> got_games = ev.play(0, /*is_evaluator*/ true);
        int opponent_moves = 0;
        for (int i = 0; i < n; ++i) {
            EXPECT(got_games[i].action_history == ref_games[i].action_history, "evaluator game %d: actions differ from the actor path", i);
            EXPECT(got_games[i].root_values == ref_games[i].root_values, "evaluator game %d: root values differ", i);
            EXPECT(got_games[i].child_visits == ref_games[i].child_visits, "evaluator game %d: stored policies differ", i);
            opponent_moves += (int)(got_games[i].action_history.size() - 1 - got_games[i].root_values.size());
        }
        EXPECT(opponent_moves > 0, "the Random opponent never moved");
    }
    // the same games once more on the same cohort: generators have advanced, histories stay well-formed
    std::vector<GameHistory> again = cohort.play(/*training_step*/ 200);
    for (int i = 0; i < n; ++i) {
        EXPECT(again[i].action_history.size() == again[i].root_values.size() + 1, "game %d: history lengths", i);
        for (size_t m = 0; m + 1 < again[i].action_history.size(); ++m) {
            const auto& legal = again[i].legal_actions[m];
            EXPECT(std::find(legal.begin(), legal.end(), again[i].action_history[m + 1]) != legal.end(), "game %d move %zu: illegal action", i, m);
            double sum = 0;
            for (double v : again[i].child_visits[m]) sum += v;
            EXPECT(std::fabs(sum - 1.0) < 1e-12, "game %d move %zu: policy sums to %g", i, m, sum);
        }
    }
}

// (7) DeviceManager with several models (two engines on the same GPU stand in for two devices): loans go to the model with the
//     fewest outstanding requests, MCTS objects spread over the models, and device 0 is left to the learner while
//     SetLearning(true) (device_manager.h:68-96, muzero.cpp:142-148)
static void device_manager_policies(const std::string& golden_dir) {
    muzero_config::MuZeroConfig config = tiny_config(4);
    DeviceManager dm;
    for (int d = 0; d < 2; ++d) {
        VPRNetModel model(config, "cuda:0", MZB_FP32_CHECK);
        model.LoadCheckpoint(golden_dir + "/checkpoint_tiny");
        dm.AddDevice(std::move(model));
    }
    EXPECT(dm.Count() == 2, "Count() = %d", dm.Count());
    {
        DeviceManager::DeviceLoan a = dm.Get(5);          // both idle: device 0
        DeviceManager::DeviceLoan b = dm.Get(3);          // device 1 has fewer outstanding requests
        DeviceManager::DeviceLoan c = dm.Get(1);          // 5 vs 3: device 1 again
        DeviceManager::DeviceLoan d = dm.Get(1, 0);       // explicit device
        EXPECT(a.device_id() == 0 && b.device_id() == 1 && c.device_id() == 1 && d.device_id() == 0, "loans went to %d %d %d %d",
               a.device_id(), b.device_id(), c.device_id(), d.device_id());
    }
    {
        DeviceManager::DeviceLoan a = dm.Get(1);          // everything returned: device 0 again
        EXPECT(a.device_id() == 0, "loans were not returned");
        dm.SetLearning(true);
        DeviceManager::DeviceLoan b = dm.Get(1), c = dm.Get(7);
        EXPECT(b.device_id() == 1 && c.device_id() == 1, "device 0 took inference work while learning");
        dm.SetLearning(false);
    }
    auto evaluator = std::make_shared<VPRNetEvaluator>(&dm, 2, 1, 2, 1);
    std::vector<std::unique_ptr<algorithm::MCTS>> mcts;
    for (int t = 0; t < 4; ++t) mcts.emplace_back(new algorithm::MCTS(config, 40 + t, evaluator));   // two per model
    std::vector<MCTSReturn> got(4);
    std::vector<std::thread> pool;
    const Observation obs(60, 0.25f);
    for (int t = 0; t < 4; ++t) pool.emplace_back([&, t] { got[t] = mcts[t]->run(obs, {0, 1, 2, 3, 4}, t % 2, false); });
    for (auto& th : pool) th.join();
    // same weights, same observation, no noise: objects on different models with the same to_play differ only through their seeds'
    // shuffles, so root values agree to rounding and every policy is a distribution
    for (int t = 0; t < 4; ++t) {
        double sum = 0;
        for (double v : got[t].emperical_policy) sum += v;
        EXPECT(std::fabs(sum - 1.0) < 1e-12 && got[t].child_actions.size() == 5, "search %d on a two-model manager is malformed", t);
    }
    dm.SetLearning(true);
    algorithm::MCTS late(config, 99, evaluator);           // must land on device 1 although it already holds two searchers
    MCTSReturn r = late.run(obs, {0, 1, 2, 3, 4}, 0, false);
    EXPECT(r.child_actions.size() == 5, "search while learning is malformed");
}

// (8) VPRNetModel::SaveCheckpoint(step) -> path (vprnet.cpp:58-63) and LoadCheckpoint(step) (vprnet.cpp:66-69) under the
//     reference's names: what one model saves another loads, and both answer identically
static void save_and_reload(const std::string& golden_dir) {
    muzero_config::MuZeroConfig config = tiny_config(2);
    char tmpl[] = "/tmp/mzb_compat_XXXXXX";
    const char* dir = mkdtemp(tmpl);
    EXPECT(dir != nullptr, "mkdtemp failed");
    if (!dir) return;
    config.path = dir;
    VPRNetModel a(config, "cuda:0", MZB_FP32_CHECK);
    a.LoadCheckpoint(golden_dir + "/checkpoint_tiny");
    const std::string saved = a.SaveCheckpoint(12);
    EXPECT(saved == std::string(dir) + "/checkpoints/checkpoint-12", "SaveCheckpoint returned %s", saved.c_str());
    VPRNetModel b(config, "cuda:0", MZB_FP32_CHECK);
    b.LoadCheckpoint(12);
    std::vector<VPRNetModel::InitialInferenceInputs> in(3);
    for (int i = 0; i < 3; ++i) { in[i].stacked_observation.assign(60, 0.f); for (int k = 0; k < 60; ++k) in[i].stacked_observation[k] = (float)((k * 7 + i * 13) % 11) / 11.f; }
    auto oa = a.InitialInference(in), ob = b.InitialInference(in);
    for (int i = 0; i < 3; ++i)
        EXPECT(oa[i].value == ob[i].value && oa[i].policy == ob[i].policy && oa[i].encoded_state == ob[i].encoded_state, "reloaded model answers differently (sample %d)", i);
}

// (9) VPRNetModel::Learn(batch) (vprnet.cpp:192-311) under the reference's name: repeated calls on one batch drive its loss down,
//     the model answers inference with the trained tensors right away, and a checkpoint saved after training reloads to the same
//     answers (the numerics of one step are checked against the unmodified reference in tests/test_learn.py)
static void learn_under_the_reference_name(const std::string& golden_dir) {
    muzero_config::MuZeroConfig config = tiny_config(2);
    char tmpl[] = "/tmp/mzb_compat_learn_XXXXXX";
    const char* dir = mkdtemp(tmpl);
    if (!dir) { EXPECT(false, "mkdtemp failed"); return; }
    config.path = dir;
    config.num_unroll_steps = 3;
    config.value_loss_weight = 0.25;
    config.network_config.learning_rate = 3e-3;
    VPRNetModel m(config, "cuda:0", MZB_FP32_CHECK);
    m.LoadCheckpoint(golden_dir + "/checkpoint_tiny");
    const int B = 8, K1 = 4, A = (int)config.action_space.size();
    Batch batch;
    batch.num_samples = B;
    std::mt19937 rng(5);
    std::uniform_real_distribution<double> u(0.0, 1.0);
    for (int b = 0; b < B; ++b) {
        batch.indices.push_back(b);
        batch.priorities.push_back(0.5 + 0.5 * u(rng));
        batch.gradient_scale.push_back(1.0 / config.num_unroll_steps);
        for (int i = 0; i < 60; ++i) batch.stacked_observations.push_back((float)u(rng));
        for (int k = 0; k < K1; ++k) {
            batch.actions.push_back((int)(rng() % A));
            batch.target_values.push_back(-4 + 8 * u(rng));
            batch.target_rewards.push_back(-2 + 4 * u(rng));
            std::vector<double> p(A);
            double sum = 0;
            for (double& x : p) { x = u(rng); sum += x; }
            for (double x : p) batch.target_policies.push_back(x / sum);
        }
    }
    std::vector<VPRNetModel::InitialInferenceInputs> in(1);
    in[0].stacked_observation.assign(batch.stacked_observations.begin(), batch.stacked_observations.begin() + 60);
    const auto before = m.InitialInference(in);
    VPRNetModel::LossInfo first = m.Learn(batch), last = first;
    EXPECT(first.errors.size() == (size_t)B && first.indices == batch.indices, "LossInfo fields");
    for (int it = 0; it < 60; ++it) last = m.Learn(batch, /*commit=*/it == 59);
    EXPECT(std::isfinite(last.total_loss) && last.total_loss < 0.8 * first.total_loss, "training did not reduce the loss: %g -> %g", first.total_loss, last.total_loss);
    const auto after = m.InitialInference(in);
    EXPECT(after[0].policy != before[0].policy, "inference still answers with the untrained weights");
    const std::string saved = m.SaveCheckpoint(3);
    VPRNetModel r(config, "cuda:0", MZB_FP32_CHECK);
    r.LoadCheckpoint(3);
    const auto again = r.InitialInference(in);
    EXPECT(again[0].policy == after[0].policy && again[0].value == after[0].value, "a checkpoint saved after training reloads differently");
    std::printf("compat_test: Learn x61 on one batch: total loss %.4f -> %.4f\n", first.total_loss, last.total_loss);
}

int main(int argc, char** argv) {
    try {
        reference_unit_tests();
        appendix_b();
        if (argc > 1) { network_paths(argv[1]); self_play_cohort(argv[1]); device_manager_policies(argv[1]); save_and_reload(argv[1]); learn_under_the_reference_name(argv[1]); }
    } catch (const std::exception& e) {
        std::printf("compat_test: %s\n", e.what());
        return 1;
    }
    if (failures) { std::printf("compat_test: %d failure(s)\n", failures); return 1; }
    std::printf("compat_test: ok\n");
    return 0;
}
