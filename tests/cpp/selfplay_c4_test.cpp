// selfplay_c4_test — the cohort drivers against the REAL reference self-play layer.
//
// tests/golden/selfplay_c4.txt was written by oracle/ref_build/ref_selfplay.cpp: 128 Connect4 games through the UNMODIFIED play_game
// (self_play.cpp:84-183), each history then re-searched by the UNMODIFIED reanalyze (self_play.cpp:186-225), and 16 evaluator games
// against the Random opponent, all over the observation-keyed synthetic evaluator.  Here the same games run through
// SelfPlayCohort / ReanalyzeCohort (include/muzero_cpp_mzb200.hpp) on an engine with MZB_EVAL_SYNTHETIC_OBS — every live game
// searched in ONE lock-step device search per move — and must come out action for action, value for value (doubles compared
// bit for bit): termination, history fields, legal sets, to_play, the search RNG stream, the actors' discrete_distribution stream,
// the opponent's uniform_int stream.
//
// The Connect4 rules below are written from the game's definition (7 columns x 6 rows, gravity, four in a row, observation =
// {player-1 stones, player-2 stones, plane of +-1 for the side to move}, reward 10 for the winning move as in the example's
// wrapper); that they agree with the reference's examples/connect4/connect4.h is itself checked: the golden file carries a hash
// of every observation the reference produced.
#include <cmath>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <map>
#include <sstream>

#include "muzero_cpp_mzb200.hpp"
#include "../../oracle/se_evaluator.h"   // test infrastructure: the observation hash of the golden file

using namespace muzero_cpp;
using namespace muzero_cpp::types;

static int failures = 0;
#define EXPECT(cond, ...) do { if (!(cond)) { if (++failures <= 20) { std::printf("FAIL %s:%d: ", __FILE__, __LINE__); std::printf(__VA_ARGS__); std::printf("\n"); } } } while (0)

class Connect4 : public AbstractGame {
public:
    Observation reset() override { std::memset(b_, 0, sizeof(b_)); mover_ = 1; return observe(); }
    StepReturn step(Action col) override {
        int row = 0;
        while (row < 6 && b_[row][col] != 0) ++row;
        b_[row][col] = mover_;
        const bool win = four(row, col);
        const bool done = win || legal_actions().empty();
        mover_ = -mover_;
        return {observe(), win ? 10.0 : 0.0, done};
    }
    Player to_play() const override { return mover_ == 1 ? 0 : 1; }
    std::vector<Action> legal_actions() const override {
        std::vector<Action> a;
        for (int c = 0; c < 7; ++c) if (b_[5][c] == 0) a.push_back(c);
        return a;
    }
    Action expert_action() override { return legal_actions()[0]; }
    Action human_to_action() override { return legal_actions()[0]; }
    std::string action_to_string(Action a) const override { return std::to_string(a); }

private:
    Observation observe() const {
        Observation o(3 * 42);
        for (int r = 0; r < 6; ++r)
            for (int c = 0; c < 7; ++c) {
                o[r * 7 + c] = b_[r][c] == 1 ? 1.f : 0.f;
                o[42 + r * 7 + c] = b_[r][c] == -1 ? 1.f : 0.f;
                o[84 + r * 7 + c] = (float)mover_;
            }
        return o;
    }
    bool four(int r, int c) const {
        static const int dr[4] = {0, 1, 1, 1}, dc[4] = {1, 0, 1, -1};
        for (int d = 0; d < 4; ++d) {
            int n = 1;
            for (int s = -1; s <= 1; s += 2)
                for (int k = 1; k < 4; ++k) {
                    const int rr = r + s * k * dr[d], cc = c + s * k * dc[d];
                    if (rr < 0 || rr >= 6 || cc < 0 || cc >= 7 || b_[rr][cc] != mover_) break;
                    ++n;
                }
            if (n >= 4) return true;
        }
        return false;
    }
    int b_[6][7];
    int mover_ = 1;
};

struct GoldenGame {
    std::vector<int> actions, to_play;
    std::vector<double> rewards, root, visits, re_root, re_visits;
    std::vector<uint32_t> legal, obs_keys;
};
struct GoldenEval { int muzero_player; double length, total_reward, mean_value; };

static double hexd(const std::string& s) { return std::strtod(s.c_str(), nullptr); }

static bool load(const std::string& path, int* sims, int* seed, int* rseed, int* eseed, std::vector<GoldenGame>* games, std::vector<GoldenEval>* evals) {
    std::ifstream in(path);
    if (!in) return false;
    std::string line;
    while (std::getline(in, line)) {
        std::istringstream ls(line);
        std::string tag, tok;
        ls >> tag;
        if (tag == "selfplay_c4") {
            std::string k; int v;
            while (ls >> k >> v) { if (k == "sims") *sims = v; if (k == "seed") *seed = v; if (k == "rseed") *rseed = v; if (k == "eseed") *eseed = v; }
        } else if (tag == "game") {
            games->emplace_back();
        } else if (tag == "eval") {
            GoldenEval e{};
            std::string k, v; int idx;
            ls >> idx;
            while (ls >> k >> v) { if (k == "muzero_player") e.muzero_player = std::atoi(v.c_str()); if (k == "length") e.length = hexd(v); if (k == "total_reward") e.total_reward = hexd(v); if (k == "mean_value") e.mean_value = hexd(v); }
            evals->push_back(e);
        } else if (!games->empty()) {
            GoldenGame& g = games->back();
            while (ls >> tok) {
                if (tag == "play_actions") g.actions.push_back(std::atoi(tok.c_str()));
                else if (tag == "play_to_play") g.to_play.push_back(std::atoi(tok.c_str()));
                else if (tag == "play_rewards") g.rewards.push_back(hexd(tok));
                else if (tag == "play_legal") g.legal.push_back((uint32_t)std::strtoul(tok.c_str(), nullptr, 10));
                else if (tag == "play_obs_keys") g.obs_keys.push_back((uint32_t)std::strtoul(tok.c_str(), nullptr, 10));
                else if (tag == "play_root_values") g.root.push_back(hexd(tok));
                else if (tag == "play_child_visits") g.visits.push_back(hexd(tok));
                else if (tag == "reanalyze_root_values") g.re_root.push_back(hexd(tok));
                else if (tag == "reanalyze_child_visits") g.re_visits.push_back(hexd(tok));
            }
        }
    }
    return !games->empty();
}

static uint32_t mask_of(const std::vector<Action>& legal) { uint32_t m = 0; for (Action a : legal) m |= 1u << a; return m; }

static muzero_config::MuZeroConfig c4_config(int sims) {
    muzero_config::MuZeroConfig c;
    c.observation_shape = {3, 6, 7};
    c.action_space = {0, 1, 2, 3, 4, 5, 6};
    c.action_channels = 1; c.num_players = 2;
    c.value_upperbound = 10; c.value_lowerbound = -10;
    c.min_reward = -10; c.max_reward = 10; c.min_value = -10; c.max_value = 10;
    c.num_simulations = sims;
    c.visit_softmax_temperature = [](int) { return 1.0; };
    return c;
}
static std::shared_ptr<mzb200::Engine> se_engine(const muzero_config::MuZeroConfig& c, int trees) {
    mzb_config pod = detail::to_pod(c, trees, MZB_FP32_CHECK, MZB_EVAL_SYNTHETIC_OBS, nullptr);
    pod.se_value_lo = -1; pod.se_value_hi = 1; pod.se_reward_lo = -0.5; pod.se_reward_hi = 0.5;   // ref_selfplay.cpp's evaluator
    return std::make_shared<mzb200::Engine>(pod, 0);
}

int main(int argc, char** argv) {
    const std::string path = argc > 1 ? std::string(argv[1]) + "/selfplay_c4.txt" : "tests/golden/selfplay_c4.txt";
    int sims = 50, seed = 0, rseed = 1000, eseed = 2000;
    std::vector<GoldenGame> gold;
    std::vector<GoldenEval> evals;
    if (!load(path, &sims, &seed, &rseed, &eseed, &gold, &evals)) { std::printf("selfplay_c4_test: cannot read %s\n", path.c_str()); return 1; }
    const int G = (int)gold.size();
    muzero_config::MuZeroConfig config = c4_config(sims);

    // ---- (1) actor games: SelfPlayCohort == play_game, game for game
    std::vector<GameHistory> played;
    {
        auto engine = se_engine(config, G);
        std::vector<Connect4> boards(G);
        std::vector<AbstractGame*> games;
        std::vector<int> seeds;
        for (int g = 0; g < G; ++g) { games.push_back(&boards[g]); seeds.push_back(seed + g); }
        SelfPlayCohort cohort(config, engine, games, seeds, seeds);
        played = cohort.play(0);
        long positions = 0;
        for (int g = 0; g < G; ++g) {
            const GameHistory& h = played[g];
            const GoldenGame& w = gold[g];
            EXPECT(h.action_history == w.actions, "game %d: actions differ (%zu vs %zu moves)", g, h.action_history.size(), w.actions.size());
            if (h.action_history != w.actions) continue;
            EXPECT(h.to_play_history == w.to_play, "game %d: to_play history differs", g);
            EXPECT(h.reward_history == w.rewards, "game %d: rewards differ", g);
            EXPECT(h.root_values == w.root, "game %d: root values differ", g);
            std::vector<double> flat;
            for (const auto& cv : h.child_visits) flat.insert(flat.end(), cv.begin(), cv.end());
            EXPECT(flat == w.visits, "game %d: stored policies differ", g);
            for (size_t i = 0; i < h.legal_actions.size(); ++i) EXPECT(mask_of(h.legal_actions[i]) == w.legal[i], "game %d step %zu: legal set differs", g, i);
            for (size_t i = 0; i < h.observation_history.size(); ++i)
                EXPECT(se_obs_key(h.observation_history[i].data(), (int)h.observation_history[i].size()) == w.obs_keys[i], "game %d step %zu: observation differs from the reference's", g, i);
            positions += (long)h.root_values.size();
        }
        std::printf("selfplay_c4_test: %d actor games, %ld searched positions, %ld lock-step searches\n", G, positions, cohort.cohort_moves());
    }

    // ---- (2) reanalyze: ReanalyzeCohort == reanalyze(), history for history (fresh engine: slot r seeded rseed + r)
    {
        auto engine = se_engine(config, G);
        std::vector<int> seeds;
        for (int g = 0; g < G; ++g) seeds.push_back(rseed + g);
        ReanalyzeCohort re(config, engine, seeds);
        std::vector<GameHistory> copy = played;
        std::vector<GameHistory*> ptrs;
        for (auto& h : copy) ptrs.push_back(&h);
        re.reanalyze(ptrs);
        for (int g = 0; g < G; ++g) {
            EXPECT(copy[g].root_values == gold[g].re_root, "history %d: reanalysed root values differ", g);
            std::vector<double> flat;
            for (const auto& cv : copy[g].child_visits) flat.insert(flat.end(), cv.begin(), cv.end());
            EXPECT(flat == gold[g].re_visits, "history %d: reanalysed policies differ", g);
        }
    }

    // ---- (3) evaluator games against the Random opponent (temperature 0, no noise, the opponent drawing from the actor's generator)
    for (int side = 0; side < 2; ++side) {
        std::vector<int> idx;
        for (int g = 0; g < (int)evals.size(); ++g) if (evals[g].muzero_player == side) idx.push_back(g);
        if (idx.empty()) continue;
        muzero_config::MuZeroConfig ec = config;
        ec.opponent_type = OpponentTypes::Random;
        ec.muzero_player = side;
        auto engine = se_engine(ec, (int)idx.size());
        std::vector<Connect4> boards(idx.size());
        std::vector<AbstractGame*> games;
        std::vector<int> seeds;
        for (size_t k = 0; k < idx.size(); ++k) { games.push_back(&boards[k]); seeds.push_back(eseed + idx[k]); }
        SelfPlayCohort cohort(ec, engine, games, seeds, seeds);
        std::vector<GameHistory> hist = cohort.play(0, true);
        for (size_t k = 0; k < idx.size(); ++k) {
            const GoldenEval& w = evals[idx[k]];
            const GameHistory& h = hist[k];
            double total = 0, mean = 0;
            for (double r : h.reward_history) total += r;
            for (double v : h.root_values) mean += v;
            mean /= (double)h.root_values.size();
            EXPECT((double)(h.action_history.size() - 1) == w.length, "evaluator game %d: length %zu vs %g", idx[k], h.action_history.size() - 1, w.length);
            EXPECT(total == w.total_reward, "evaluator game %d: total reward %g vs %g", idx[k], total, w.total_reward);
            EXPECT(std::fabs(mean - w.mean_value) <= 1e-12, "evaluator game %d: mean root value %.17g vs %.17g", idx[k], mean, w.mean_value);
        }
    }
    // ---- (4) replay: the played histories through buffer::PrioritizedReplayBuffer, event for event as the UNMODIFIED reference
    // buffer handled them (tests/golden/replay_c4.txt) — here the generator is a real std::mt19937 behind the C-ABI callback
    {
        const std::string rpath = argc > 1 ? std::string(argv[1]) + "/replay_c4.txt" : "tests/golden/replay_c4.txt";
        std::ifstream in(rpath);
        std::string line;
        if (!std::getline(in, line)) { std::printf("selfplay_c4_test: cannot read %s\n", rpath.c_str()); return 1; }
        std::map<std::string, std::string> head;
        { std::istringstream ls(line); std::string tag, k, v; ls >> tag; while (ls >> k >> v) head[k] = v; }
        muzero_config::MuZeroConfig rc = config;
        rc.stacked_observations = std::atoi(head["stack"].c_str());
        rc.td_steps = std::atoi(head["td"].c_str()); rc.num_unroll_steps = std::atoi(head["unroll"].c_str());
        rc.per_alpha = hexd(head["alpha"]); rc.per_beta = hexd(head["beta"]); rc.per_beta_increment = hexd(head["beta_increment"]);
        rc.per_epsilon = hexd(head["epsilon"]); rc.discount = hexd(head["discount"]); rc.min_sample_size = std::atoi(head["min_sample_size"].c_str());
        rc.action_representation_initial = [](Action a) { return Observation(42, (float)((double)a / 7)); };
        auto engine = se_engine(rc, 8);
        buffer::PrioritizedReplayBuffer rb(rc, std::atoi(head["capacity"].c_str()), "replay", engine);
        std::mt19937 rng(std::atoi(head["seed"].c_str()));
        Batch last;
        int batches = 0;
        auto ints = [](std::istringstream& ls) { std::vector<int> v; int x; while (ls >> x) v.push_back(x); return v; };
        auto dbls = [](std::istringstream& ls) { std::vector<double> v; std::string t; while (ls >> t) v.push_back(hexd(t)); return v; };
        while (std::getline(in, line)) {
            std::istringstream ls(line);
            std::string tag;
            ls >> tag;
            if (tag == "save") {
                int g, size, can; std::string k;
                ls >> g >> k >> size >> k >> can;
                rb.save_game_history(played[g]);
                EXPECT(rb.size() == size && (int)rb.can_sample() == can, "replay: size after game %d is %d, reference %d", g, rb.size(), size);
            } else if (tag == "sample") {
                int n; ls >> n;
                last = rb.sample(rng, n);
                ++batches;
            } else if (tag == "indices") { EXPECT(last.indices == ints(ls), "replay batch %d: indices differ", batches);
            } else if (tag == "actions") { EXPECT(last.actions == ints(ls), "replay batch %d: actions differ", batches);
            } else if (tag == "priorities") { EXPECT(last.priorities == dbls(ls), "replay batch %d: importance weights differ", batches);
            } else if (tag == "target_values") { EXPECT(last.target_values == dbls(ls), "replay batch %d: target values differ", batches);
            } else if (tag == "target_rewards") { EXPECT(last.target_rewards == dbls(ls), "replay batch %d: target rewards differ", batches);
            } else if (tag == "target_policies") { EXPECT(last.target_policies == dbls(ls), "replay batch %d: target policies differ", batches);
            } else if (tag == "gradient_scale") { EXPECT(last.gradient_scale == dbls(ls), "replay batch %d: gradient scale differs", batches);
            } else if (tag == "stacked_observations") {
                std::vector<float> want; double x;
                while (ls >> x) want.push_back((float)(x / 7.0));   // printed as value * 7: 0, 7 (a stone) or the action of an action plane
                EXPECT(last.stacked_observations == want, "replay batch %d: stacked observations differ", batches);
            } else if (tag == "update") {
                std::vector<double> e = dbls(ls);
                rb.update_history_priorities(last.indices, e);
            } else if (tag == "sample_game") {
                int id; size_t len;
                ls >> id >> len;
                auto sg = rb.sample_game(rng);
                EXPECT(std::get<0>(sg) == id && std::get<1>(sg).root_values.size() == len, "replay: sample_game %d (%zu positions), reference %d (%zu)",
                       std::get<0>(sg), std::get<1>(sg).root_values.size(), id, len);
                if (len > 0) EXPECT(std::get<1>(sg).action_history == played[id - 1].action_history, "replay: sample_game returned another history");
            }
        }
        EXPECT(batches == 16, "replay: %d batches compared", batches);
        std::printf("selfplay_c4_test: %d replay batches\n", batches);
    }
    if (failures) { std::printf("selfplay_c4_test: %d failure(s)\n", failures); return 1; }
    std::printf("selfplay_c4_test: ok\n");
    return 0;
}
