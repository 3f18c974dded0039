// mzb200_facade.hpp — C++ host facade over the C ABI (mzb200.h) with the reference's own call shapes, so that
// self_play.cpp / reanalyze keep calling `mcts.run(observation, legal_actions, to_play, add_noise)` and
// `model.InitialInference(batch)` while the work happens in libmzb200.so.  Header-only, C++17, no torch, no abseil.
//
//   reference type / call (file:line)                          facade
//   ---------------------------------------------------------  -----------------------------------------------------------
//   types::MCTSReturn (muzero-cpp/types.h:70-76)                mzb200::MCTSReturn (same five fields, same order)
//   algorithm::MCTS::MCTS(config, seed, evaluator)             mzb200::MCTS(cohort, seed): takes a tree slot, seeds its mt19937
//     (mcts.cpp:17-31)
//   MCTS::run(obs, legal_actions, to_play, add_noise)          mzb200::MCTS::run: same arguments, same return; blocks until the
//     (mcts.cpp:34-87; callers self_play.cpp:130,208)            cohort of concurrent actors has been searched in lock step
//   VPRNetEvaluator request queue + 1 ms deadline               mzb200::Cohort: the last actor to arrive (or the deadline) launches
//     (vprnet_evaluator.cpp:105-178)                              ONE mzb_search_batch_slots call for every parked actor
//   VPRNetModel::{Initial,Recurrent}Inference (vprnet.cpp:77-189)  mzb200::Model::{InitialInference,RecurrentInference}
//   DeviceManager::AddDevice / Get (device_manager.h:36-80)     one mzb200::Engine per GPU; no loans (the model is resident)
//
// Error behaviour mirrors the reference (assert / std::exit, no error codes of its own): a failing C-ABI call prints
// mzb_last_error() and aborts, unless MZB200_FACADE_THROW is defined, in which case std::runtime_error is thrown.
#pragma once
#include <chrono>
#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <memory>
#include <mutex>
#include <stdexcept>
#include <string>
#include <vector>

#include "mzb200.h"

namespace mzb200 {

using Action = int;                        // types.h:18
using Player = int;                        // types.h:20
using Observation = std::vector<float>;    // types.h:22

struct MCTSReturn {                        // types.h:70-76
    double root_value = 0;
    int max_tree_depth = 0;
    std::vector<double> emperical_policy;
    std::vector<double> children_relative_visit;
    std::vector<Action> child_actions;
};

inline void check(int rc, const char* what) {
    if (rc == MZB_OK) return;
    std::string msg = std::string(what) + ": " + mzb_last_error();
#ifdef MZB200_FACADE_THROW
    throw std::runtime_error(msg);
#else
    std::fprintf(stderr, "mzb200: %s\n", msg.c_str());
    std::abort();
#endif
}

inline void fail(const std::string& msg) {
#ifdef MZB200_FACADE_THROW
    throw std::runtime_error(msg);
#else
    std::fprintf(stderr, "mzb200: %s\n", msg.c_str());
    std::abort();
#endif
}

// legal_actions -> the engine's bit mask.  The engine creates the root's children in ascending action order; the reference
// creates them in the order of the caller's vector (mcts.cpp:163-166), and that order feeds std::shuffle and the first-strict-
// maximum tie rule.  The two agree exactly when the vector is strictly ascending — what every AbstractGame::legal_actions() of
// the reference returns (connect4.h, tictactoe.h, the ALE wrapper) — so anything else is refused instead of silently searched
// in another order; out-of-range actions would shift by a negative or too large amount.
inline uint32_t legal_mask_of(const std::vector<Action>& legal_actions, int num_actions) {
    uint32_t mask = 0;
    Action prev = -1;
    for (Action a : legal_actions) {
        if (a < 0 || a >= num_actions || a >= 32) fail("legal action " + std::to_string(a) + " outside the action space [0, " + std::to_string(num_actions) + ")");
        if (a <= prev) fail("legal_actions must be strictly ascending (the engine expands the root's children in action order)");
        prev = a;
        mask |= 1u << a;
    }
    return mask;
}

// One engine = one GPU = one resident model + max_trees tree slots (replaces VPRNetModel + DeviceManager + the MCTS
// objects' storage for that device).
class Engine {
public:
    Engine(const mzb_config& config, int device) : cfg_(config) { check(mzb_engine_create(&cfg_, device, &e_), "mzb_engine_create"); }
    ~Engine() { mzb_engine_destroy(e_); }
    Engine(const Engine&) = delete;
    Engine& operator=(const Engine&) = delete;
    mzb_engine* handle() const { return e_; }
    const mzb_config& config() const { return cfg_; }
    // VPRNetModel::LoadCheckpoint (vprnet.cpp:66-74): the module's named fp32 tensors
    void LoadWeights(const std::vector<mzb_tensor_desc>& tensors) { check(mzb_load_weights(e_, tensors.data(), (int)tensors.size()), "mzb_load_weights"); }
    mzb_shapes shapes() const { mzb_shapes s; check(mzb_get_shapes(e_, &s), "mzb_get_shapes"); return s; }
    // An engine is driven by one host thread at a time (mzb200.h): callers that share it (a Cohort's searches next to
    // VPRNetEvaluator batches) serialise on this mutex.
    std::mutex& call_mutex() { return call_mu_; }

private:
    mzb_config cfg_;
    mzb_engine* e_ = nullptr;
    std::mutex call_mu_;
};

// VPRNetModel's batch inference API (vprnet.h:64-110) on host vectors.
struct InferenceOutputs {   // vprnet.h:64-69
    double value = 0, reward = 0;
    std::vector<double> policy;
    Observation encoded_state;
};
class Model {
public:
    explicit Model(std::shared_ptr<Engine> engine) : eng_(std::move(engine)), sh_(eng_->shapes()) {}
    std::vector<InferenceOutputs> InitialInference(const std::vector<Observation>& stacked_observations) {
        const int B = (int)stacked_observations.size(), A = eng_->config().num_actions;
        std::vector<float> obs((size_t)B * sh_.initial_flat_size);
        for (int b = 0; b < B; ++b) {
            if (stacked_observations[b].size() != (size_t)sh_.initial_flat_size)
                fail("InitialInference: observation " + std::to_string(b) + " has " + std::to_string(stacked_observations[b].size()) + " values, the network input has " + std::to_string(sh_.initial_flat_size));
            std::copy(stacked_observations[b].begin(), stacked_observations[b].end(), obs.begin() + (size_t)b * sh_.initial_flat_size);
        }
        std::vector<double> v(B), r(B), p((size_t)B * A);
        std::vector<float> h((size_t)B * sh_.recurrent_flat_size);
        std::lock_guard<std::mutex> call(eng_->call_mutex());
        check(mzb_initial_inference(eng_->handle(), obs.data(), B, v.data(), r.data(), p.data(), h.data()), "mzb_initial_inference");
        return unpack(B, A, v, r, p, h);
    }
    std::vector<InferenceOutputs> RecurrentInference(const std::vector<Observation>& hidden_states, const std::vector<Action>& actions) {
        const int B = (int)hidden_states.size(), A = eng_->config().num_actions;
        std::vector<float> hin((size_t)B * sh_.recurrent_flat_size);
        if (actions.size() != hidden_states.size()) fail("RecurrentInference: " + std::to_string(hidden_states.size()) + " hidden states, " + std::to_string(actions.size()) + " actions");
        for (int b = 0; b < B; ++b) {
            if (hidden_states[b].size() != (size_t)sh_.recurrent_flat_size)
                fail("RecurrentInference: hidden state " + std::to_string(b) + " has " + std::to_string(hidden_states[b].size()) + " values, expected " + std::to_string(sh_.recurrent_flat_size));
            std::copy(hidden_states[b].begin(), hidden_states[b].end(), hin.begin() + (size_t)b * sh_.recurrent_flat_size);
        }
        std::vector<int32_t> act(actions.begin(), actions.end());
        std::vector<double> v(B), r(B), p((size_t)B * A);
        std::vector<float> h((size_t)B * sh_.recurrent_flat_size);
        std::lock_guard<std::mutex> call(eng_->call_mutex());
        check(mzb_recurrent_inference(eng_->handle(), hin.data(), act.data(), B, v.data(), r.data(), p.data(), h.data()), "mzb_recurrent_inference");
        return unpack(B, A, v, r, p, h);
    }

private:
    std::vector<InferenceOutputs> unpack(int B, int A, const std::vector<double>& v, const std::vector<double>& r,
                                         const std::vector<double>& p, const std::vector<float>& h) const {
        std::vector<InferenceOutputs> out(B);
        for (int b = 0; b < B; ++b) {
            out[b].value = v[b]; out[b].reward = r[b];
            out[b].policy.assign(p.begin() + (size_t)b * A, p.begin() + (size_t)(b + 1) * A);
            out[b].encoded_state.assign(h.begin() + (size_t)b * sh_.recurrent_flat_size, h.begin() + (size_t)(b + 1) * sh_.recurrent_flat_size);
        }
        return out;
    }
    std::shared_ptr<Engine> eng_;
    mzb_shapes sh_;
};

// The batching point of the blocking one-tree API.  Actors (one thread per game, as in the reference: self_play.cpp:228-238)
// call MCTS::run; the call parks here until every registered actor has a request pending or `deadline` has passed since
// the first one arrived (the reference's evaluator uses a 1 ms deadline for the same purpose, vprnet_evaluator.cpp:118);
// the actor that completes the cohort runs ONE lock-step search for all of them and wakes the others.
class Cohort {
public:
    Cohort(std::shared_ptr<Engine> engine, std::chrono::microseconds deadline = std::chrono::microseconds(1000))
        : eng_(std::move(engine)), deadline_(deadline), used_(eng_->config().max_trees, 0) {
        width_ = eng_->config().evaluator == MZB_EVAL_SYNTHETIC ? 1 : (size_t)eng_->shapes().initial_flat_size;
    }

    std::shared_ptr<Engine> engine() const { return eng_; }
    // MCTS objects currently holding a tree slot of this engine
    int members() { std::lock_guard<std::mutex> lk(mu_); return members_; }

    // MCTS::MCTS: take a free tree slot and seed its std::mt19937 (mcts.cpp:26)
    int acquire_slot(int seed) {
        std::lock_guard<std::mutex> lk(mu_);
        for (int s = 0; s < (int)used_.size(); ++s)
            if (!used_[s]) {
                used_[s] = 1; ++members_;
                const int32_t sd = seed;
                std::lock_guard<std::mutex> call(eng_->call_mutex());
                check(mzb_search_seed(eng_->handle(), s, 1, &sd), "mzb_search_seed");
                return s;
            }
        std::fprintf(stderr, "mzb200: more MCTS objects than engine tree slots (max_trees = %d)\n", (int)used_.size());
        std::abort();
    }
    void release_slot(int slot) {
        std::lock_guard<std::mutex> lk(mu_);
        used_[slot] = 0; --members_;
        cv_.notify_all();   // the parked cohort may now be complete
    }

    MCTSReturn search(int slot, const Observation& obs, const std::vector<Action>& legal_actions, Player to_play, bool add_noise) {
        Request rq;
        rq.slot = slot; rq.obs = &obs; rq.to_play = to_play; rq.noise = add_noise ? 1 : 0;
        // the reference hands the observation to a torch tensor of the network's input shape and fails there on a wrong size
        // (vprnet.cpp:84-90); here it would be read or written past its end, so it is checked before it is queued
        if (obs.size() != width_) fail("MCTS::run: observation has " + std::to_string(obs.size()) + " values, the network input has " + std::to_string(width_));
        rq.mask = legal_mask_of(legal_actions, eng_->config().num_actions);
        std::unique_lock<std::mutex> lk(mu_);
        cv_.wait(lk, [&] { return !running_; });        // a search is in flight: join the next cohort
        if (pending_.empty()) first_arrival_ = std::chrono::steady_clock::now();
        pending_.push_back(&rq);
        cv_.notify_all();
        while (!rq.done) {
            if (!rq.taken && !running_ &&
                ((int)pending_.size() >= members_ || std::chrono::steady_clock::now() - first_arrival_ >= deadline_)) {
                launch(lk);                              // this actor completes the cohort: it runs the search for everyone
            } else if (!rq.taken && !running_) {
                cv_.wait_until(lk, first_arrival_ + deadline_);
            } else {
                cv_.wait(lk);
            }
        }
        return std::move(rq.result);
    }

private:
    struct Request {
        int slot = 0; const Observation* obs = nullptr; uint32_t mask = 0; int32_t to_play = 0; uint8_t noise = 0;
        bool taken = false, done = false; MCTSReturn result;
    };
    // called with the lock held by whichever thread completes the cohort; runs the search unlocked
    void launch(std::unique_lock<std::mutex>& lk) {
        std::vector<Request*> batch;
        batch.swap(pending_);
        for (Request* r : batch) r->taken = true;
        running_ = true;
        lk.unlock();
        const int n = (int)batch.size(), A = eng_->config().num_actions;
        const size_t width = width_;   // every queued observation has exactly this size (checked in search())
        std::vector<float> obs((size_t)n * width);
        std::vector<int32_t> slots(n), to_play(n), n_children(n), depth(n), actions((size_t)n * A);
        std::vector<uint32_t> masks(n);
        std::vector<uint8_t> noise(n);
        std::vector<double> root(n), rel((size_t)n * A), pol((size_t)n * A);
        for (int i = 0; i < n; ++i) {
            std::copy(batch[i]->obs->begin(), batch[i]->obs->end(), obs.begin() + (size_t)i * width);
            slots[i] = batch[i]->slot; to_play[i] = batch[i]->to_play; masks[i] = batch[i]->mask; noise[i] = batch[i]->noise;
        }
        mzb_search_result out{};
        out.root_value = root.data(); out.max_tree_depth = depth.data(); out.n_children = n_children.data();
        out.child_actions = actions.data(); out.children_relative_visit = rel.data(); out.emperical_policy = pol.data();
        {
            std::lock_guard<std::mutex> call(eng_->call_mutex());
            check(mzb_search_batch_slots(eng_->handle(), obs.data(), slots.data(), masks.data(), to_play.data(), noise.data(), n, &out),
                  "mzb_search_batch_slots");
        }
        lk.lock();
        for (int i = 0; i < n; ++i) {
            MCTSReturn& r = batch[i]->result;
            const int L = n_children[i];
            r.root_value = root[i]; r.max_tree_depth = depth[i];
            r.emperical_policy.assign(pol.begin() + (size_t)i * A, pol.begin() + (size_t)(i + 1) * A);
            r.children_relative_visit.assign(rel.begin() + (size_t)i * A, rel.begin() + (size_t)i * A + L);
            r.child_actions.assign(actions.begin() + (size_t)i * A, actions.begin() + (size_t)i * A + L);
            batch[i]->done = true;
        }
        running_ = false;
        cv_.notify_all();
    }

    std::shared_ptr<Engine> eng_;
    std::chrono::microseconds deadline_;
    size_t width_ = 0;
    std::mutex mu_;
    std::condition_variable cv_;
    std::vector<char> used_;
    int members_ = 0;
    bool running_ = false;
    std::vector<Request*> pending_;
    std::chrono::steady_clock::time_point first_arrival_;
};

// algorithm::MCTS (mcts.h:113-200): thread-confined like the reference's (one per actor).
class MCTS {
public:
    MCTS(std::shared_ptr<Cohort> cohort, int seed) : cohort_(std::move(cohort)), slot_(cohort_->acquire_slot(seed)) {}
    ~MCTS() { cohort_->release_slot(slot_); }
    MCTS(const MCTS&) = delete;
    MCTS& operator=(const MCTS&) = delete;
    MCTSReturn run(const Observation& stacked_observation, const std::vector<Action>& legal_actions, Player to_play,
                   bool add_exploration_noise) {
        return cohort_->search(slot_, stacked_observation, legal_actions, to_play, add_exploration_noise);
    }
    int slot() const { return slot_; }

private:
    std::shared_ptr<Cohort> cohort_;
    int slot_;
};

}  // namespace mzb200
