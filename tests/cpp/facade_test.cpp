// facade_test.cpp — the reference's actor pattern (one thread per game calling MCTS::run, self_play.cpp:228-238;
// checks in the spirit of tests/mcts_test.cpp) through mzb200_facade.hpp -> libmzb200.so.
// Uses the synthetic evaluator, so no weights are needed; compares the parked-actor path against one direct batched call.
#define MZB200_FACADE_THROW
#include <cmath>
#include <cstdio>
#include <cstring>
#include <thread>

#include "mzb200_facade.hpp"

static mzb_config make_config(int trees) {
    mzb_config c;
    std::memset(&c, 0, sizeof(c));
    c.observation_c = 1; c.observation_h = 2; c.observation_w = 2;
    c.num_actions = 7; c.action_channels = 1; c.num_players = 2; c.stacked_observations = 0;
    c.value_upperbound = 10; c.value_lowerbound = -10;
    c.min_reward = -10; c.max_reward = 10; c.min_value = -10; c.max_value = 10; c.use_contractive_mapping = 1;
    c.num_simulations = 50; c.discount = 0.997; c.dirichlet_alpha = 0.3; c.dirichlet_epsilon = 0.25;
    c.pb_c_base = 19652; c.pb_c_init = 1.25; c.seed = 0;
    c.resnet_channels = 4; c.normalize_hidden_states = 1;
    c.max_trees = trees; c.precision = MZB_FP32_CHECK; c.evaluator = MZB_EVAL_SYNTHETIC;
    c.se_value_lo = -1; c.se_value_hi = 1; c.se_reward_lo = 0; c.se_reward_hi = 1;
    return c;
}

int main() {
    const int T = 8, moves = 3, A = 7;
    try {
        // (a) direct batched calls: slot t seeded 100 + t
        std::vector<std::vector<mzb200::MCTSReturn>> want(moves, std::vector<mzb200::MCTSReturn>(T));
        {
            mzb200::Engine eng(make_config(T), 0);
            std::vector<int32_t> seeds(T);
            for (int t = 0; t < T; ++t) seeds[t] = 100 + t;
            mzb200::check(mzb_search_seed(eng.handle(), 0, T, seeds.data()), "seed");
            for (int m = 0; m < moves; ++m) {
                std::vector<float> obs(T);
                std::vector<uint32_t> mask(T, 0x7f);
                std::vector<int32_t> tp(T), nch(T), depth(T), act(T * A);
                std::vector<uint8_t> noise(T, 1);
                std::vector<double> root(T), rel(T * A), pol(T * A);
                for (int t = 0; t < T; ++t) { obs[t] = (float)(1 + t + 17 * m); tp[t] = (t + m) % 2; }
                mzb_search_result out{};
                out.root_value = root.data(); out.max_tree_depth = depth.data(); out.n_children = nch.data();
                out.child_actions = act.data(); out.children_relative_visit = rel.data(); out.emperical_policy = pol.data();
                mzb200::check(mzb_search_batch(eng.handle(), obs.data(), mask.data(), tp.data(), noise.data(), T, &out), "batch");
                for (int t = 0; t < T; ++t) {
                    want[m][t].root_value = root[t]; want[m][t].max_tree_depth = depth[t];
                    want[m][t].child_actions.assign(act.begin() + t * A, act.begin() + t * A + nch[t]);
                    want[m][t].children_relative_visit.assign(rel.begin() + t * A, rel.begin() + t * A + nch[t]);
                    want[m][t].emperical_policy.assign(pol.begin() + t * A, pol.begin() + (t + 1) * A);
                }
            }
        }
        // (b) T actor threads, each with its own MCTS object, blocking run() calls
        auto eng = std::make_shared<mzb200::Engine>(make_config(T), 0);
        auto cohort = std::make_shared<mzb200::Cohort>(eng, std::chrono::microseconds(200000));
        std::vector<std::unique_ptr<mzb200::MCTS>> mcts(T);
        for (int t = 0; t < T; ++t) mcts[t].reset(new mzb200::MCTS(cohort, 100 + t));   // slots are handed out in order
        int bad = 0;
        std::vector<std::thread> actors;
        std::mutex mu;
        for (int t = 0; t < T; ++t)
            actors.emplace_back([&, t] {
                for (int m = 0; m < moves; ++m) {
                    mzb200::Observation obs = {(float)(1 + t + 17 * m)};
                    mzb200::MCTSReturn r = mcts[t]->run(obs, {0, 1, 2, 3, 4, 5, 6}, (t + m) % 2, true);
                    const mzb200::MCTSReturn& w = want[m][t];
                    double sum = 0;
                    for (double x : r.children_relative_visit) sum += x;
                    const bool ok = r.root_value == w.root_value && r.max_tree_depth == w.max_tree_depth && r.child_actions == w.child_actions &&
                                    r.children_relative_visit == w.children_relative_visit && r.emperical_policy == w.emperical_policy &&
                                    std::fabs(sum - 1.0) < 1e-12;
                    if (!ok) { std::lock_guard<std::mutex> lk(mu); ++bad; }
                }
            });
        for (auto& th : actors) th.join();
        // (c) ragged cohort: half of the actors retire, the rest keep searching (deadline / membership path)
        for (int t = T / 2; t < T; ++t) mcts[t].reset();
        std::vector<std::thread> rest;
        for (int t = 0; t < T / 2; ++t)
            rest.emplace_back([&, t] {
                mzb200::Observation obs = {(float)(5 + t)};
                mzb200::MCTSReturn r = mcts[t]->run(obs, {1, 3, 5}, 0, false);
                if (r.child_actions.size() != 3) { std::lock_guard<std::mutex> lk(mu); ++bad; }
            });
        for (auto& th : rest) th.join();
        std::printf("facade_test: %s (%d mismatches)\n", bad ? "FAILED" : "ok", bad);
        return bad ? 1 : 0;
    } catch (const std::exception& e) {
        std::printf("facade_test: exception: %s\n", e.what());
        return 2;
    }
}
