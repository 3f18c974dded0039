// ref_selfplay — TEST INFRASTRUCTURE.  Drives the UNMODIFIED reference's self-play layer (muzero-cpp/self_play.cpp, compiled where
// it lies by oracle/ref_build/Makefile) and dumps what it produces, as golden fixtures for the cohort drivers of the B200 engine
// (include/muzero_cpp_mzb200.hpp: SelfPlayCohort, ReanalyzeCohort).  Never linked into the product.
//
//   play_game (self_play.cpp:84-183) and reanalyze (self_play.cpp:186-225) have external linkage but no declaration in self_play.h:
//   they are declared here with the reference's signatures and called directly.
//
//   * actor games:  for game g: Connect4 (examples/connect4/connect4.h behind an AbstractGame wrapper that scales the reward by 10,
//     like examples/connect4/muzero_connect4.cpp:32-36), MCTS(config, seed + g, evaluator), std::mt19937(seed + g) — the pairing of
//     self_play_actor (self_play.cpp:228-238) — one play_game call, the GameHistory popped from the trajectory queue.
//   * reanalyze:    the history of game g goes ALONE into a PrioritizedReplayBuffer; reanalyze() runs on MCTS(config, rseed + g) and
//     re-searches every stored position once.  reanalyze() only returns on a stop request: the evaluator requests it when the
//     first search of the SECOND pass begins, and the function leaves at its next position without storing that pass
//     (self_play.cpp:205,219) — so exactly one full pass reaches the buffer.
//   * evaluator games (is_evaluator = true, opponent Random): temperature 0, no noise, the opponent's moves drawn from the SAME
//     generator as select_action would use (self_play.cpp:41-56,128-143); only SharedStats sees their outcome.
//
//   * replay (replay_out=FILE): every played history also goes into ONE PrioritizedReplayBuffer (replay_buffer.cpp, unmodified) of a
//     capacity small enough that old games are evicted; after every `replay_every` games one sample(rng, batch) is taken and dumped
//     in full (types::Batch), its priorities are updated from synthetic errors, and one sample_game(rng) is drawn — all from one
//     std::mt19937(replay_seed).  Golden for mzb_replay_* (csrc/replay.cu).
//
// Evaluator = "SE" keyed by the whole observation (oracle/se_evaluator.h, se_obs_key) — MZB_EVAL_SYNTHETIC_OBS on the engine side.
#include <cinttypes>
#include <cstdio>
#include <map>
#include <memory>
#include <random>
#include <string>
#include <vector>

#include "examples/connect4/connect4.h"
#include "muzero-cpp/abstract_game.h"
#include "muzero-cpp/config.h"
#include "muzero-cpp/mcts.h"
#include "muzero-cpp/queue.h"
#include "muzero-cpp/replay_buffer.h"
#include "muzero-cpp/self_play.h"
#include "muzero-cpp/shared_stats.h"
#include "muzero-cpp/types.h"
#include "muzero-cpp/util.h"

#include "../se_evaluator.h"

using namespace muzero_cpp;
using namespace muzero_cpp::types;
using muzero_cpp::muzero_config::MuZeroConfig;

namespace muzero_cpp {
void play_game(const muzero_config::MuZeroConfig& config, algorithm::MCTS& mcts, AbstractGame& game,
               const std::shared_ptr<SharedStats>& shared_stats, std::mt19937& rng,
               ThreadedQueue<GameHistory>* trajectory_queue, bool is_evaluator, bool render, util::StopToken* stop);
void reanalyze(const muzero_config::MuZeroConfig& config, algorithm::MCTS& mcts,
               std::shared_ptr<buffer::PrioritizedReplayBuffer> reanalyze_buffer, std::shared_ptr<SharedStats> shared_stats,
               std::mt19937& rng, util::StopToken* stop);
}  // namespace muzero_cpp

class Connect4Game : public AbstractGame {   // the wrapper of examples/connect4/muzero_connect4.cpp:17-98
public:
    explicit Connect4Game(int seed) : env_(seed) {}
    Observation reset() override { return env_.reset(); }
    StepReturn step(Action a) override { StepReturn r = env_.step(a); r.reward *= 10; return r; }
    Player to_play() const override { return env_.to_play(); }
    std::vector<Action> legal_actions() const override { return env_.legal_actions(); }
    Action expert_action() override { return env_.expert_action(); }
    Action human_to_action() override { return env_.legal_actions()[0]; }
    std::string action_to_string(Action a) const override { return std::to_string(a); }
private:
    Connect4 env_;
};

class SEObsEvaluator : public Evaluator {
public:
    SEObsEvaluator(int na) : na_(na) {}
    VPRNetModel::InferenceOutputs InitialInference(const Observation& obs) override {
        ++initial_calls;
        if (stop_at > 0 && initial_calls >= stop_at && stop) stop->stop();
        return make(se_initial_cid(se_obs_key(obs.data(), (int)obs.size())));
    }
    VPRNetModel::InferenceOutputs RecurrentInference(const Observation& hidden, Action action) override {
        return make(se_child_cid((uint32_t)hidden[0], (uint32_t)action));
    }
    long initial_calls = 0, stop_at = 0;
    util::StopToken* stop = nullptr;
private:
    VPRNetModel::InferenceOutputs make(uint32_t cid) {
        VPRNetModel::InferenceOutputs o;
        o.value = se_value(cid, -1.0, 1.0);
        o.reward = se_reward(cid, -0.5, 0.5);
        o.policy.resize(na_);
        se_policy(cid, na_, o.policy.data());
        o.encoded_state = {(float)cid};
        return o;
    }
    int na_;
};

static Observation encode_action(Action a) { return Observation(42, (double)a / 7); }

static MuZeroConfig connect4_config(int sims, const std::string& path) {
    MuZeroConfig c;
    c.seed = 0;
    c.path = path;
    c.observation_shape = Connect4::obs_shape();
    c.action_space = Connect4::action_space();
    c.action_channels = 1; c.num_players = 2;
    c.value_upperbound = 10; c.value_lowerbound = -10;
    c.min_reward = -10; c.max_reward = 10; c.min_value = -10; c.max_value = 10;
    c.num_simulations = sims;
    c.min_sample_size = 1;
    c.action_representation_initial = encode_action;
    c.action_representation_recurrent = encode_action;
    c.visit_softmax_temperature = [](int) { return 1.0; };
    return c;
}

static uint32_t mask_of(const std::vector<Action>& legal) { uint32_t m = 0; for (Action a : legal) m |= 1u << a; return m; }

static void dump_history(FILE* f, const char* tag, const GameHistory& h) {
    fprintf(f, "%s_actions", tag); for (Action a : h.action_history) fprintf(f, " %d", a); fprintf(f, "\n");
    fprintf(f, "%s_rewards", tag); for (double r : h.reward_history) fprintf(f, " %a", r); fprintf(f, "\n");
    fprintf(f, "%s_to_play", tag); for (Player p : h.to_play_history) fprintf(f, " %d", p); fprintf(f, "\n");
    fprintf(f, "%s_legal", tag); for (auto& l : h.legal_actions) fprintf(f, " %u", mask_of(l)); fprintf(f, "\n");
    fprintf(f, "%s_obs_keys", tag); for (auto& o : h.observation_history) fprintf(f, " %u", se_obs_key(o.data(), (int)o.size())); fprintf(f, "\n");
}
static void dump_search(FILE* f, const char* tag, const GameHistory& h) {
    fprintf(f, "%s_root_values", tag); for (double v : h.root_values) fprintf(f, " %a", v); fprintf(f, "\n");
    fprintf(f, "%s_child_visits", tag);
    for (auto& cv : h.child_visits) for (double v : cv) fprintf(f, " %a", v);
    fprintf(f, "\n");
}

static void dump_batch(FILE* f, const Batch& b) {
    fprintf(f, "indices"); for (int v : b.indices) fprintf(f, " %d", v); fprintf(f, "\n");
    fprintf(f, "priorities"); for (double v : b.priorities) fprintf(f, " %a", v); fprintf(f, "\n");
    fprintf(f, "actions"); for (int v : b.actions) fprintf(f, " %d", v); fprintf(f, "\n");
    fprintf(f, "target_values"); for (double v : b.target_values) fprintf(f, " %a", v); fprintf(f, "\n");
    fprintf(f, "target_rewards"); for (double v : b.target_rewards) fprintf(f, " %a", v); fprintf(f, "\n");
    fprintf(f, "target_policies"); for (double v : b.target_policies) fprintf(f, " %a", v); fprintf(f, "\n");
    fprintf(f, "gradient_scale"); for (double v : b.gradient_scale) fprintf(f, " %a", v); fprintf(f, "\n");
    // observations are 0/1 planes and a/7 action planes: print the distinct values compactly (value index per cell)
    fprintf(f, "stacked_observations"); for (float v : b.stacked_observations) fprintf(f, " %g", (double)v * 7); fprintf(f, "\n");
}

int main(int argc, char** argv) {
    std::map<std::string, std::string> kv;
    for (int i = 1; i < argc; ++i) { std::string s(argv[i]); auto p = s.find('='); if (p != std::string::npos) kv[s.substr(0, p)] = s.substr(p + 1); }
    auto geti = [&](const char* k, int d) { return kv.count(k) ? std::stoi(kv[k]) : d; };
    const int games = geti("games", 8), sims = geti("sims", 50), seed = geti("seed", 0), rseed = geti("rseed", 1000);
    const int eval_games = geti("eval_games", 0), eseed = geti("eseed", 2000);
    const std::string path = kv.count("path") ? kv["path"] : "/tmp/mzb_ref_selfplay";
    FILE* f = kv.count("out") ? fopen(kv["out"].c_str(), "w") : stdout;
    MuZeroConfig config = connect4_config(sims, path);
    fprintf(f, "selfplay_c4 games %d sims %d seed %d rseed %d eval_games %d eseed %d\n", games, sims, seed, rseed, eval_games, eseed);
    auto stats = std::make_shared<SharedStats>();
    // replay golden
    FILE* rf = kv.count("replay_out") ? fopen(kv["replay_out"].c_str(), "w") : nullptr;
    const int r_cap = geti("replay_capacity", 1000), r_batch = geti("replay_batch", 16), r_stack = geti("replay_stack", 2);
    const int r_every = geti("replay_every", 8), r_seed = geti("replay_seed", 4242);
    MuZeroConfig rbc = config;
    rbc.path = path + "/replay";
    rbc.stacked_observations = r_stack;
    rbc.td_steps = geti("replay_td", 5); rbc.num_unroll_steps = geti("replay_unroll", 5);
    rbc.per_alpha = 0.6; rbc.per_beta = 0.4; rbc.per_beta_increment = 0.05; rbc.per_epsilon = 0.01; rbc.discount = 0.997;
    rbc.min_sample_size = 32;
    std::unique_ptr<buffer::PrioritizedReplayBuffer> rb;
    std::mt19937 replay_rng(r_seed);
    if (rf) {
        rb.reset(new buffer::PrioritizedReplayBuffer(rbc, r_cap, "replay"));
        fprintf(rf, "replay_c4 games %d capacity %d batch %d stack %d every %d seed %d td %d unroll %d alpha %a beta %a beta_increment %a epsilon %a discount %a min_sample_size %d\n",
                games, r_cap, r_batch, r_stack, r_every, r_seed, rbc.td_steps, rbc.num_unroll_steps, rbc.per_alpha, rbc.per_beta,
                rbc.per_beta_increment, rbc.per_epsilon, rbc.discount, rbc.min_sample_size);
    }
    for (int g = 0; g < games; ++g) {
        auto ev = std::make_shared<SEObsEvaluator>(7);
        util::StopToken stop;
        Connect4Game game(seed + g);
        algorithm::MCTS mcts(config, seed + g, ev);
        std::mt19937 rng(seed + g);
        ThreadedQueue<GameHistory> queue(4);
        play_game(config, mcts, game, stats, rng, &queue, false, false, &stop);
        absl::optional<GameHistory> got = queue.Pop();
        if (!got) { fprintf(stderr, "no history for game %d\n", g); return 1; }
        GameHistory h = *got;
        fprintf(f, "game %d positions %zu\n", g, h.root_values.size());
        dump_history(f, "play", h);
        dump_search(f, "play", h);
        if (rf) {
            rb->save_game_history(h);
            fprintf(rf, "save %d size %d can_sample %d\n", g, rb->size(), (int)rb->can_sample());
            if ((g + 1) % r_every == 0 && rb->can_sample()) {
                Batch b = rb->sample(replay_rng, r_batch);
                fprintf(rf, "sample %d\n", b.num_samples);
                dump_batch(rf, b);
                std::vector<double> errors;
                for (int idx : b.indices) errors.push_back(((unsigned)idx * 2654435761u % 1000u) / 100.0 - 3.0);
                fprintf(rf, "update"); for (double e : errors) fprintf(rf, " %a", e); fprintf(rf, "\n");
                rb->update_history_priorities(b.indices, errors);
                auto sg = rb->sample_game(replay_rng);
                fprintf(rf, "sample_game %d %zu\n", std::get<0>(sg), std::get<1>(sg).root_values.size());
            }
        }
        // reanalyze: this history alone in a buffer, one full pass (see the header comment)
        MuZeroConfig rc = config;
        rc.path = path + "/re" + std::to_string(g);
        auto buffer = std::make_shared<buffer::PrioritizedReplayBuffer>(rc, 1000, "reanalyze");
        buffer->save_game_history(h);
        auto rev = std::make_shared<SEObsEvaluator>(7);
        util::StopToken rstop;
        rev->stop = &rstop; rev->stop_at = (long)h.root_values.size() + 1;
        algorithm::MCTS rmcts(rc, rseed + g, rev);
        std::mt19937 rrng(rseed + g);
        reanalyze(rc, rmcts, buffer, stats, rrng, &rstop);
        std::mt19937 pick(0);
        GameHistory after = std::get<1>(buffer->sample_game(pick));
        dump_search(f, "reanalyze", after);
    }
    for (int g = 0; g < eval_games; ++g) {   // evaluator games against the Random opponent, muzero_player = g % 2
        MuZeroConfig ec = config;
        ec.opponent_type = OpponentTypes::Random;
        ec.muzero_player = g % 2;
        auto ev = std::make_shared<SEObsEvaluator>(7);
        util::StopToken stop;
        Connect4Game game(eseed + g);
        algorithm::MCTS mcts(ec, eseed + g, ev);
        std::mt19937 rng(eseed + g);
        auto es = std::make_shared<SharedStats>();
        play_game(ec, mcts, game, es, rng, nullptr, true, false, &stop);
        SharedStats::Stats st = es->get_all_stats();
        fprintf(f, "eval %d muzero_player %d length %a total_reward %a mean_value %a\n", g, ec.muzero_player, st.evaluator_episode_length,
                st.evaluator_total_reward, st.evaluator_mean_value);
    }
    if (rf) fclose(rf);
    if (f != stdout) fclose(f);
    return 0;
}
