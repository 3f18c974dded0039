// replay.cu — replay sampling and training targets: PrioritizedReplayBuffer::{save_game_history, sample,
// update_history_priorities, sample_game} (muzero-cpp/replay_buffer.cpp:51-141) over SumTree (replay_buffer.h:28-209), and
// GameHistory::{compute_target_value, make_target, get_stacked_observations} (types.cpp:16-95).
//
// Split by what each side is good at:
//   * HOST: the sum tree itself (2 * capacity - 1 doubles), the (step, history) of every leaf and the lengths of the histories — a few
//     bytes per stored position.  One sample() walks it B times (log2(capacity) comparisons each) and draws from the CALLER's
//     generator through libstdc++'s own std::uniform_real_distribution / std::uniform_int_distribution, in the reference's order
//     (one real per sample, then one int per absorbing unroll step of that sample), so the random stream is consumed identically.
//   * DEVICE (HBM): everything with volume — the histories' observation frames, actions, rewards, to_play, root values and visit
//     distributions, stored once when a game is saved.  One launch of replay_batch_kernel assembles the whole batch from them:
//     n-step value targets, reward / policy / action targets over the unroll, absorbing steps, gradient scale and the stacked
//     observation of every sample (a gather of frames and action planes straight into the batch, HBM-bound; for ALE-shaped
//     frames that is 479 KB per sample that never crosses PCIe).  f64 arithmetic in the reference's operation order, built without
//     FMA contraction; std::pow(discount, i) comes from a host table (glibc), like the search's log table.
// The assembled batch is returned in host buffers (mzb_batch layout) and stays resident for a learner on the same GPU.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <deque>
#include <random>
#include <unordered_map>

#include "engine.h"

namespace mzb {
namespace replay_detail {

// the caller's generator behind the interface std::mt19937 shows to libstdc++'s distributions (32-bit draws)
struct CallerRng {
    typedef uint_fast32_t result_type;
    mzb_rng_fn fn; void* user;
    static constexpr result_type min() { return 0; }
    static constexpr result_type max() { return 0xffffffffUL; }
    result_type operator()() { return (result_type)fn(user); }
};

struct HistMeta {
    int length = 0;            // root_values.size()
    int n_actions_hist = 0;    // action_history.size() = length + 1
    long row0 = 0;             // first row of this history in the per-position device arrays
    int refs = 0;              // leaves of the tree that still point at it (SumTree::hist_count_)
};

}  // namespace replay_detail
}  // namespace mzb

using namespace mzb;
using namespace mzb::replay_detail;

struct mzb_replay {
    mzb_engine* e = nullptr;
    // PrioritizedReplayBuffer's scalars (replay_buffer.cpp:15-32)
    double alpha = 1, beta = 1, epsilon = 0.01, beta_increment = 0.001, discount = 0.997;
    int min_sample_size = 1, td_steps = 10, num_unroll_steps = 5, capacity = 0;
    // SumTree (replay_buffer.h:28-209)
    std::vector<double> tree;
    std::vector<std::pair<int, int>> data;     // leaf -> (step, history id)
    int num_entries = 0, position = 0, map_counter = 0;
    std::unordered_map<int, HistMeta> hist;
    std::deque<int> fifo;                      // stored histories in insertion (= eviction) order
    // device store: per-position rows in a ring, [rows][...]
    long rows = 0, row_head = 0;               // next free row; histories are contiguous, evicted first-in first-out
    int frame = 0, A = 0;
    DevBuf d_frames, d_actions, d_rewards, d_to_play, d_root, d_visits, d_pow, d_planes;
    // per-sample descriptors and outputs of the last batch
    DevBuf d_desc, d_abs_actions, d_tv, d_tr, d_tp, d_act, d_gs, d_obs;
    int batch_cap = 0;
};

namespace mzb { namespace replay_detail {

void tree_update(mzb_replay* r, int index, double priority) {   // SumTree::update
    double change = priority - r->tree[index];
    r->tree[index] = priority;
    while (index != 0) {
        index = (index - 1) / 2;
        r->tree[index] += change;
    }
}
int tree_retrieve(const mzb_replay* r, double value) {          // SumTree::retrieve, iteratively
    int index = 0;
    while (true) {
        const int left = 2 * index + 1, right = left + 1;
        if (left >= (int)r->tree.size()) return index;
        if (value <= r->tree[left]) index = left;
        else { value -= r->tree[left]; index = right; }
    }
}
// GameHistory::compute_target_value (types.cpp:45-70) on host arrays (initial priorities of a saved game)
double host_target_value(const mzb_game_history* h, int step, int td_steps, double discount) {
    const int bootstrap = step + td_steps;
    double value = 0;
    if (bootstrap < h->length) {
        value = h->root_values[bootstrap] * std::pow(discount, td_steps);
        if (h->to_play[step] != h->to_play[bootstrap]) value *= -1;
    }
    const int start = step + 1, end = std::min(bootstrap + 1, h->length + 1);
    for (int i = 0; i < end - start; ++i) {
        const bool same = h->to_play[step] == h->to_play[step + i];
        const double reward = same ? h->rewards[start + i] : -h->rewards[start + i];
        value += reward * std::pow(discount, i);
    }
    return value;
}

// ---------------------------------------------------------------- batch assembly on the device

struct SampleDesc { long row0; int length, step, abs_off; };   // abs_off: first of this sample's pre-drawn absorbing actions

// One block per sample.  make_target (types.cpp:73-95) + the gradient scale and stacked observation of sample() (replay_buffer.cpp:
// 108-121).  Thread 0 walks the n-step return (a dependent f64 chain of <= td_steps terms); the other threads fill rewards, actions
// and policies meanwhile; then the whole block gathers the stacked observation plane by plane (coalesced copies).
__global__ void replay_batch_kernel(const SampleDesc* __restrict__ desc, const int* __restrict__ abs_actions, const float* __restrict__ frames,
                                    const int* __restrict__ actions, const double* __restrict__ rewards, const int* __restrict__ to_play,
                                    const double* __restrict__ root, const double* __restrict__ visits, const double* __restrict__ powt,
                                    const float* __restrict__ planes, int td_steps, int K, int A, int c, int hw, int ach, int S,
                                    double* __restrict__ tv, double* __restrict__ tr, double* __restrict__ tp, int* __restrict__ act,
                                    double* __restrict__ gs, float* __restrict__ obs) {
    const int b = blockIdx.x, K1 = K + 1;
    const SampleDesc d = desc[b];
    const int L = d.length, step = d.step;
    __shared__ double s_value;
    if (threadIdx.x == 0) {   // compute_target_value(step, td_steps, discount) (types.cpp:45-70) — the SAME value for every unroll step
        const int bootstrap = step + td_steps;
        double value = 0;
        if (bootstrap < L) {
            value = root[d.row0 + bootstrap] * powt[td_steps];
            if (to_play[d.row0 + step] != to_play[d.row0 + bootstrap]) value *= -1;
        }
        const int start = step + 1, end = min(bootstrap + 1, L + 1);
        for (int i = 0; i < end - start; ++i) {
            const bool same = to_play[d.row0 + step] == to_play[d.row0 + step + i];
            const double reward = same ? rewards[d.row0 + start + i] : -rewards[d.row0 + start + i];
            value += reward * powt[i];
        }
        s_value = value;
        gs[b] = 1.0 / (double)min(K, L + 1 - step);   // 1 / min(num_unroll_steps, action_history.size() - step)
    }
    // absorbing steps are past the end of the game: uniform policy, zero value / reward, a random action (drawn on the host from
    // the caller's generator, in order); one step past the end still has its reward and action (types.cpp:84-93)
    for (int k = threadIdx.x; k < K1; k += blockDim.x) {
        const int i = step + k;
        const bool non_abs_end = i <= L;
        tr[(long)b * K1 + k] = non_abs_end ? rewards[d.row0 + i] : 0.0;
        int n_abs_before = 0;                       // absorbing steps among step .. i - 1 (they are the last ones of the unroll)
        if (!non_abs_end) n_abs_before = i - max(L + 1, step);
        act[(long)b * K1 + k] = non_abs_end ? actions[d.row0 + i] : abs_actions[d.abs_off + n_abs_before];
    }
    for (int j = threadIdx.x; j < K1 * A; j += blockDim.x) {
        const int k = j / A, a = j - k * A, i = step + k;
        tp[((long)b * K1 + k) * A + a] = i < L ? visits[(d.row0 + i) * A + a] : 1.0 / A;
    }
    __syncthreads();
    for (int k = threadIdx.x; k < K1; k += blockDim.x) tv[(long)b * K1 + k] = (step + k < L) ? s_value : 0.0;
    // get_stacked_observations(step, S, ...) (types.cpp:16-42): current frame, then for past = step-1 .. step-S: frame + action planes
    const int planes_out = c + S * (c + ach);
    float* out = obs + (long)b * planes_out * hw;
    for (int p = 0; p < planes_out; ++p) {
        const float* src = nullptr;
        if (p < c) src = frames + ((d.row0 + step) * c + p) * (long)hw;
        else {
            const int r = p - c, per = c + ach, past = step - 1 - r / per, j = r % per;
            if (past >= 0) {
                if (j < c) src = frames + ((d.row0 + past) * c + j) * (long)hw;
                else src = planes + ((long)actions[d.row0 + past + 1] * ach + (j - c)) * hw;
            }
        }
        for (int i = threadIdx.x; i < hw; i += blockDim.x) out[(long)p * hw + i] = src ? src[i] : 0.f;
    }
}

int ensure_batch(mzb_replay* r, int B) {
    if (B <= r->batch_cap) return MZB_OK;
    const mzb_config& c = r->e->cfg;
    const int K1 = r->num_unroll_steps + 1, flat = r->e->dims.initial_flat();
    MZB_CUDA(r->d_desc.alloc((size_t)B * sizeof(SampleDesc)));
    MZB_CUDA(r->d_abs_actions.alloc((size_t)B * K1 * sizeof(int)));
    MZB_CUDA(r->d_tv.alloc((size_t)B * K1 * sizeof(double)));
    MZB_CUDA(r->d_tr.alloc((size_t)B * K1 * sizeof(double)));
    MZB_CUDA(r->d_tp.alloc((size_t)B * K1 * c.num_actions * sizeof(double)));
    MZB_CUDA(r->d_act.alloc((size_t)B * K1 * sizeof(int)));
    MZB_CUDA(r->d_gs.alloc((size_t)B * sizeof(double)));
    MZB_CUDA(r->d_obs.alloc((size_t)B * flat * sizeof(float)));
    r->batch_cap = B;
    return MZB_OK;
}

} }  // namespace mzb::replay_detail

extern "C" {

int mzb_replay_create(mzb_engine* e, const mzb_replay_config* rc, mzb_replay** out) {
    MZB_REQUIRE(e && rc && out, "null argument");
    MZB_REQUIRE(rc->capacity >= 1 && rc->td_steps >= 0 && rc->num_unroll_steps >= 0, "bad replay configuration");
    MZB_CUDA(cudaSetDevice(e->device));
    std::unique_ptr<mzb_replay> r(new mzb_replay());
    r->e = e;
    r->alpha = rc->per_alpha; r->beta = rc->per_beta; r->epsilon = rc->per_epsilon; r->beta_increment = rc->per_beta_increment;
    r->discount = e->cfg.discount; r->min_sample_size = rc->min_sample_size; r->td_steps = rc->td_steps;
    r->num_unroll_steps = rc->num_unroll_steps; r->capacity = rc->capacity;
    r->tree.assign((size_t)2 * rc->capacity - 1, 0.0);
    r->data.assign(rc->capacity, {0, 0});
    const mzb_config& c = e->cfg;
    r->frame = c.observation_c * c.observation_h * c.observation_w;
    r->A = c.num_actions;
    r->rows = 2L * rc->capacity + 65536;
    MZB_CUDA(r->d_frames.alloc((size_t)r->rows * r->frame * sizeof(float)));
    MZB_CUDA(r->d_actions.alloc((size_t)r->rows * sizeof(int)));
    MZB_CUDA(r->d_rewards.alloc((size_t)r->rows * sizeof(double)));
    MZB_CUDA(r->d_to_play.alloc((size_t)r->rows * sizeof(int)));
    MZB_CUDA(r->d_root.alloc((size_t)r->rows * sizeof(double)));
    MZB_CUDA(r->d_visits.alloc((size_t)r->rows * r->A * sizeof(double)));
    std::vector<double> powt(rc->td_steps + 1);
    for (int i = 0; i <= rc->td_steps; ++i) powt[i] = std::pow(r->discount, i);   // glibc pow, as the reference evaluates it
    MZB_CUDA(r->d_pow.alloc(powt.size() * sizeof(double)));
    MZB_CUDA(cudaMemcpy(r->d_pow.p, powt.data(), powt.size() * sizeof(double), cudaMemcpyHostToDevice));
    // action_representation_initial per action at observation resolution; NULL = the examples' constant plane a / A
    const int hw = c.observation_h * c.observation_w, ach = c.action_channels;
    std::vector<float> planes((size_t)r->A * std::max(1, ach) * hw, 0.f);
    for (int a = 0; a < r->A; ++a)
        for (int i = 0; i < ach * hw; ++i)
            planes[(size_t)a * ach * hw + i] = rc->action_planes_initial ? rc->action_planes_initial[(size_t)a * ach * hw + i] : (float)((double)a / r->A);
    MZB_CUDA(r->d_planes.alloc(planes.size() * sizeof(float)));
    MZB_CUDA(cudaMemcpy(r->d_planes.p, planes.data(), planes.size() * sizeof(float), cudaMemcpyHostToDevice));
    *out = r.release();
    return MZB_OK;
}

void mzb_replay_destroy(mzb_replay* r) {
    if (!r) return;
    cudaSetDevice(r->e->device);
    delete r;
}

int mzb_replay_size(const mzb_replay* r) { return r ? r->num_entries : 0; }
int mzb_replay_can_sample(const mzb_replay* r) { return r && r->num_entries >= r->min_sample_size; }
int mzb_replay_num_histories(const mzb_replay* r) { return r ? r->map_counter : 0; }
double mzb_replay_total_priority(const mzb_replay* r) { return r ? r->tree[0] : 0.0; }
// histories leave in the order they came (the leaves are overwritten first in, first out): every id below this one is gone
int mzb_replay_oldest_history(const mzb_replay* r) { return !r ? 0 : r->fifo.empty() ? r->map_counter + 1 : r->fifo.front(); }

// save_game_history (replay_buffer.cpp:51-66) + SumTree::add (replay_buffer.h:86-109)
int mzb_replay_save_game_history(mzb_replay* r, const mzb_game_history* h) {
    MZB_REQUIRE(r && h, "null argument");
    const int L = h->length;
    MZB_REQUIRE(L >= 1 && L < 32768, "history length %d outside [1, 32768)", L);
    MZB_REQUIRE(L <= r->capacity, "a history of %d positions does not fit a buffer of %d (the reference would overwrite it while adding it)", L, r->capacity);
    MZB_REQUIRE(h->observations && h->actions && h->rewards && h->to_play && h->root_values && h->child_visits, "null history field");
    MZB_CUDA(cudaSetDevice(r->e->device));
    // initial priorities: |root value - n-step return| (Appendix G), error_to_priorities (replay_buffer.cpp:45-49)
    std::vector<double> pri(L);
    for (int i = 0; i < L; ++i) pri[i] = std::pow(std::abs(h->root_values[i] - host_target_value(h, i, r->td_steps, r->discount)) + r->epsilon, r->alpha);
    // rows in the device ring (contiguous; first in, first out like the tree's leaves)
    const long need = L + 1;
    if (r->row_head + need > r->rows) r->row_head = 0;
    if (!r->fifo.empty()) {
        const HistMeta& oldest = r->hist[r->fifo.front()];
        const bool clash = oldest.row0 >= r->row_head && oldest.row0 < r->row_head + need;
        MZB_REQUIRE(!clash, "replay store: device ring exhausted (capacity %d)", r->capacity);
    }
    const long row0 = r->row_head;
    r->row_head += need;
    cudaStream_t s = r->e->stream;
    MZB_CUDA(cudaMemcpyAsync(r->d_frames.as<float>() + row0 * r->frame, h->observations, (size_t)need * r->frame * sizeof(float), cudaMemcpyHostToDevice, s));
    MZB_CUDA(cudaMemcpyAsync(r->d_actions.as<int>() + row0, h->actions, (size_t)need * sizeof(int), cudaMemcpyHostToDevice, s));
    MZB_CUDA(cudaMemcpyAsync(r->d_rewards.as<double>() + row0, h->rewards, (size_t)need * sizeof(double), cudaMemcpyHostToDevice, s));
    MZB_CUDA(cudaMemcpyAsync(r->d_to_play.as<int>() + row0, h->to_play, (size_t)need * sizeof(int), cudaMemcpyHostToDevice, s));
    MZB_CUDA(cudaMemcpyAsync(r->d_root.as<double>() + row0, h->root_values, (size_t)L * sizeof(double), cudaMemcpyHostToDevice, s));
    MZB_CUDA(cudaMemcpyAsync(r->d_visits.as<double>() + row0 * r->A, h->child_visits, (size_t)L * r->A * sizeof(double), cudaMemcpyHostToDevice, s));
    MZB_CUDA(cudaStreamSynchronize(s));
    // SumTree::add
    const int id = ++r->map_counter;
    HistMeta m;
    m.length = L; m.n_actions_hist = L + 1; m.row0 = row0;
    r->hist[id] = m;
    r->fifo.push_back(id);
    for (int step = 0; step < L; ++step) {
        const int index = r->position + r->capacity - 1;
        const int old = r->data[r->position].second;
        if (old > 0) {
            auto it = r->hist.find(old);
            if (it != r->hist.end() && --it->second.refs == 0) {
                r->num_entries -= it->second.length;
                r->hist.erase(it);
                while (!r->fifo.empty() && !r->hist.count(r->fifo.front())) r->fifo.pop_front();
            }
        }
        r->data[r->position] = {step, id};
        ++r->hist[id].refs;
        tree_update(r, index, pri[step]);
        r->position = (r->position + 1) % r->capacity;
        ++r->num_entries;
    }
    return MZB_OK;
}

// update_history_priorities (replay_buffer.cpp:69-77)
int mzb_replay_update_priorities(mzb_replay* r, const int32_t* indices, const double* errors, int n) {
    MZB_REQUIRE(r && indices && errors && n >= 0, "null argument");
    for (int i = 0; i < n; ++i) {
        MZB_REQUIRE(indices[i] >= r->capacity - 1 && indices[i] < 2 * r->capacity - 1, "index %d is not a leaf of the tree", indices[i]);
        tree_update(r, indices[i], std::pow(std::abs(errors[i]) + r->epsilon, r->alpha));
    }
    return MZB_OK;
}

// reanalyze's write-back of a re-searched history (update_game_history, replay_buffer.cpp:138-141): new root values / policies
int mzb_replay_update_game_history(mzb_replay* r, int history_id, const double* root_values, const double* child_visits, int length) {
    MZB_REQUIRE(r && root_values && child_visits, "null argument");
    auto it = r->hist.find(history_id);
    if (it == r->hist.end()) return MZB_OK;   // the buffer has lost this history meanwhile: nothing to do (as the reference)
    MZB_REQUIRE(length == it->second.length, "history %d holds %d positions, not %d", history_id, it->second.length, length);
    MZB_CUDA(cudaSetDevice(r->e->device));
    MZB_CUDA(cudaMemcpy(r->d_root.as<double>() + it->second.row0, root_values, (size_t)length * sizeof(double), cudaMemcpyHostToDevice));
    MZB_CUDA(cudaMemcpy(r->d_visits.as<double>() + it->second.row0 * r->A, child_visits, (size_t)length * r->A * sizeof(double), cudaMemcpyHostToDevice));
    return MZB_OK;
}

// sample_game (replay_buffer.cpp:80-85): a history id drawn uniformly from 1 .. histories added so far, and its length (0 if lost)
int mzb_replay_sample_game(mzb_replay* r, mzb_rng_fn next, void* user, int32_t* history_id, int32_t* length) {
    MZB_REQUIRE(r && next && history_id && length, "null argument");
    MZB_REQUIRE(r->map_counter >= 1, "empty replay buffer");
    CallerRng rng{next, user};
    std::uniform_int_distribution<> d(1, r->map_counter);
    *history_id = d(rng);
    auto it = r->hist.find(*history_id);
    *length = it == r->hist.end() ? 0 : it->second.length;
    return MZB_OK;
}

// sample (replay_buffer.cpp:88-135)
int mzb_replay_sample(mzb_replay* r, mzb_rng_fn next, void* user, int batch_size, mzb_batch_out* out) {
    MZB_REQUIRE(r && next && out && batch_size >= 1, "null argument");
    MZB_REQUIRE(r->num_entries >= 1, "empty replay buffer");
    MZB_REQUIRE(out->priorities && out->indices && out->actions && out->stacked_observations && out->target_rewards && out->target_values &&
                out->target_policies && out->gradient_scale, "null output field");
    MZB_CUDA(cudaSetDevice(r->e->device));
    int rc = ensure_batch(r, batch_size);
    if (rc) return rc;
    const int B = batch_size, K = r->num_unroll_steps, K1 = K + 1, A = r->A;
    CallerRng rng{next, user};
    r->beta = std::min(1.0, r->beta + r->beta_increment);
    const double segment = r->tree[0] / B;
    std::vector<SampleDesc> desc(B);
    std::vector<int> abs_actions;
    std::uniform_int_distribution<int> action_dist(0, A - 1);
    for (int i = 0; i < B; ++i) {
        std::uniform_real_distribution<double> u(segment * i, segment * (i + 1));
        const double value = u(rng);
        const int leaf = tree_retrieve(r, value);
        const int slot = leaf - r->capacity + 1;
        const int step = r->data[slot].first, id = r->data[slot].second;
        auto it = r->hist.find(id);
        MZB_REQUIRE(it != r->hist.end(), "sampled an empty leaf (total priority %g)", r->tree[0]);
        out->indices[i] = leaf;
        out->priorities[i] = r->tree[leaf];
        desc[i].row0 = it->second.row0; desc[i].length = it->second.length; desc[i].step = step; desc[i].abs_off = (int)abs_actions.size();
        // make_target draws one action per absorbing step, in unroll order, from the same generator (types.cpp:93)
        for (int k = step; k < step + K1; ++k)
            if (k > it->second.length) abs_actions.push_back(action_dist(rng));
    }
    // importance-sampling weights (replay_buffer.cpp:124-131)
    double mx = 0;
    for (int i = 0; i < B; ++i) {
        out->priorities[i] = std::pow(r->num_entries * out->priorities[i] / r->tree[0], -r->beta);
        mx = i == 0 ? out->priorities[i] : std::max(mx, out->priorities[i]);
    }
    for (int i = 0; i < B; ++i) out->priorities[i] /= mx;
    // the batch itself: one launch over the resident histories
    cudaStream_t s = r->e->stream;
    const mzb_config& c = r->e->cfg;
    const int flat = r->e->dims.initial_flat(), hw = c.observation_h * c.observation_w;
    abs_actions.push_back(0);
    MZB_CUDA(cudaMemcpyAsync(r->d_desc.p, desc.data(), (size_t)B * sizeof(SampleDesc), cudaMemcpyHostToDevice, s));
    MZB_CUDA(cudaMemcpyAsync(r->d_abs_actions.p, abs_actions.data(), abs_actions.size() * sizeof(int), cudaMemcpyHostToDevice, s));
    replay_batch_kernel<<<B, 128, 0, s>>>(r->d_desc.as<SampleDesc>(), r->d_abs_actions.as<int>(), r->d_frames.as<float>(), r->d_actions.as<int>(),
                                         r->d_rewards.as<double>(), r->d_to_play.as<int>(), r->d_root.as<double>(), r->d_visits.as<double>(),
                                         r->d_pow.as<double>(), r->d_planes.as<float>(), r->td_steps, K, A, c.observation_c, hw, c.action_channels,
                                         c.stacked_observations, r->d_tv.as<double>(), r->d_tr.as<double>(), r->d_tp.as<double>(), r->d_act.as<int>(),
                                         r->d_gs.as<double>(), r->d_obs.as<float>());
    MZB_CUDA(cudaGetLastError());
    MZB_CUDA(cudaMemcpyAsync(out->target_values, r->d_tv.p, (size_t)B * K1 * sizeof(double), cudaMemcpyDeviceToHost, s));
    MZB_CUDA(cudaMemcpyAsync(out->target_rewards, r->d_tr.p, (size_t)B * K1 * sizeof(double), cudaMemcpyDeviceToHost, s));
    MZB_CUDA(cudaMemcpyAsync(out->target_policies, r->d_tp.p, (size_t)B * K1 * A * sizeof(double), cudaMemcpyDeviceToHost, s));
    MZB_CUDA(cudaMemcpyAsync(out->actions, r->d_act.p, (size_t)B * K1 * sizeof(int), cudaMemcpyDeviceToHost, s));
    MZB_CUDA(cudaMemcpyAsync(out->gradient_scale, r->d_gs.p, (size_t)B * sizeof(double), cudaMemcpyDeviceToHost, s));
    MZB_CUDA(cudaMemcpyAsync(out->stacked_observations, r->d_obs.p, (size_t)B * flat * sizeof(float), cudaMemcpyDeviceToHost, s));
    MZB_CUDA(cudaStreamSynchronize(s));
    out->num_samples = B;
    return MZB_OK;
}

}  // extern "C"
