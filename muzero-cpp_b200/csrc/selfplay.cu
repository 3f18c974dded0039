// selfplay.cu — the callers' side of the search: per-slot observation history kept ON THE DEVICE and stacked
// there (GameHistory::get_stacked_observations, types.cpp:16-42), and the actors' action selection (select_action,
// self_play.cpp:21-39).  With the history resident, a move uploads ONE new frame per game instead of the whole stacked observation
// (ALE shape: 36.9 KB instead of 479 KB per game and move) and the stacked batch is built by an HBM-bound gather kernel straight
// into the search's input buffer.
//
// Device layout per tree slot: a ring of K = stacked_observations + 1 frames [K][c*h*w] fp32, the action that led to each frame,
// and the number of frames pushed since the last reset (count).  Frame t of the game sits in ring entry t % K.
#include <algorithm>
#include <cmath>
#include <random>

#include "engine.h"

namespace mzb {
int search_with_device_obs_slots(mzb_engine* e, const int32_t* slots, const uint32_t* legal_mask, const int32_t* to_play,
                                 const uint8_t* add_noise, int n, mzb_search_result* out);

struct ObsRing {
    DevBuf frames, actions, count, planes, d_slots, d_push_actions, d_push_frames;
    int K = 0, c = 0, hw = 0, ach = 0, S = 0, A = 0;
    long frame_elems = 0;
    std::vector<int> h_count;            // host mirror of count (validation only)
    std::vector<std::mt19937> actor;     // std::mt19937 rng(config.seed + actor_num) (self_play.cpp:233)
    std::vector<uint8_t> actor_seeded;
};

void obs_ring_free(ObsRing* r) { delete r; }

namespace {

// one block per (game, output plane): a plane is h*w contiguous floats
__global__ void stack_obs_kernel(const float* __restrict__ frames, const int* __restrict__ actions, const int* __restrict__ count,
                                 const float* __restrict__ planes, const int* __restrict__ slots, float* __restrict__ out,
                                 int K, int c, int hw, int ach, int S, long out_stride) {
    const int game = blockIdx.y, plane = blockIdx.x;
    const int slot = slots[game];
    const int step = count[slot] - 1;                 // index of the current observation (get_stacked_observations(-1, ...))
    const float* src = nullptr;                       // nullptr = zero plane (ran off the trajectory, types.cpp:37-40)
    if (plane < c) {
        src = frames + ((long)slot * K + step % K) * ((long)c * hw) + (long)plane * hw;
    } else {
        const int r = plane - c, per = c + ach;
        const int past = step - 1 - r / per, j = r % per;   // types.cpp:29: past_index = step-1, step-2, ...
        if (past >= 0) {
            if (j < c) src = frames + ((long)slot * K + past % K) * ((long)c * hw) + (long)j * hw;
            else src = planes + ((long)actions[(long)slot * K + (past + 1) % K] * ach + (j - c)) * hw;   // action_history[past+1]
        }
    }
    float* dst = out + (long)game * out_stride + (long)plane * hw;
    for (int i = threadIdx.x; i < hw; i += blockDim.x) dst[i] = src ? src[i] : 0.f;
}

__global__ void ring_push_kernel(float* frames, int* actions, int* count, const int* slots, const float* new_frames, const int* new_actions,
                                 int K, long frame_elems) {
    const int game = blockIdx.y, slot = slots[game];
    const int t = count[slot];                        // index of the frame being appended
    float* dst = frames + ((long)slot * K + t % K) * frame_elems;
    const float* src = new_frames + (long)game * frame_elems;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < frame_elems; i += (long)gridDim.x * blockDim.x) dst[i] = src[i];
    if (blockIdx.x == 0 && threadIdx.x == 0) actions[(long)slot * K + t % K] = t == 0 ? 0 : new_actions[game];   // action_history[0] = 0 (self_play.cpp:91)
}
__global__ void ring_count_kernel(int* count, const int* slots, int n, int reset) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) count[slots[i]] = reset ? 0 : count[slots[i]] + 1;
}

int ring_check(mzb_engine* e, const int32_t* slots, int n, bool need_frames) {
    MZB_REQUIRE(e && slots, "null argument");
    MZB_REQUIRE(e->ring, "no observation ring (mzb_obs_ring_create)");
    MZB_REQUIRE(n >= 1 && n <= e->cfg.max_trees, "n %d outside [1, max_trees=%d]", n, e->cfg.max_trees);
    std::vector<uint8_t> seen(e->cfg.max_trees, 0);
    for (int i = 0; i < n; ++i) {
        MZB_REQUIRE(slots[i] >= 0 && slots[i] < e->cfg.max_trees, "row %d: slot %d outside [0, %d)", i, slots[i], e->cfg.max_trees);
        MZB_REQUIRE(!seen[slots[i]], "slot %d appears twice", slots[i]);
        seen[slots[i]] = 1;
        if (need_frames) MZB_REQUIRE(e->ring->h_count[slots[i]] > 0, "slot %d has no observation yet (mzb_obs_ring_push)", slots[i]);
    }
    return MZB_OK;
}

int ring_stack_into(mzb_engine* e, const int32_t* slots, int n, float* d_out, cudaStream_t s) {
    ObsRing& r = *e->ring;
    MZB_CUDA(cudaMemcpyAsync(r.d_slots.p, slots, (size_t)n * sizeof(int), cudaMemcpyHostToDevice, s));
    const int planes_out = r.c + r.S * (r.c + r.ach);
    dim3 grid(planes_out, n);
    stack_obs_kernel<<<grid, 128, 0, s>>>(r.frames.as<float>(), r.actions.as<int>(), r.count.as<int>(), r.planes.as<float>(), r.d_slots.as<int>(),
                                         d_out, r.K, r.c, r.hw, r.ach, r.S, (long)planes_out * r.hw);
    MZB_CUDA(cudaGetLastError());
    return MZB_OK;
}

}  // namespace
}  // namespace mzb

using namespace mzb;

extern "C" {

int mzb_obs_ring_create(mzb_engine* e, const float* action_planes_initial) {
    MZB_REQUIRE(e, "null argument");
    MZB_REQUIRE(e->cfg.evaluator == MZB_EVAL_NETWORK || e->cfg.evaluator == MZB_EVAL_SYNTHETIC_OBS,
                "the observation ring feeds evaluators that take whole observation rows (NETWORK, SYNTHETIC_OBS)");
    MZB_CUDA(cudaSetDevice(e->device));
    const mzb_config& c = e->cfg;
    std::unique_ptr<ObsRing, ObsRingDeleter> r(new ObsRing());
    r->S = c.stacked_observations; r->K = r->S + 1; r->c = c.observation_c; r->hw = c.observation_h * c.observation_w;
    r->ach = c.action_channels; r->A = c.num_actions;
    r->frame_elems = (long)r->c * r->hw;
    MZB_REQUIRE(r->c + r->S * (r->c + r->ach) == e->dims.cin, "stacked planes %d != network input channels %d", r->c + r->S * (r->c + r->ach), e->dims.cin);
    const size_t T = c.max_trees;
    MZB_CUDA(r->frames.alloc(T * r->K * r->frame_elems * sizeof(float)));
    MZB_CUDA(r->actions.alloc(T * r->K * sizeof(int)));
    MZB_CUDA(r->count.alloc(T * sizeof(int)));
    MZB_CUDA(cudaMemset(r->count.p, 0, T * sizeof(int)));
    MZB_CUDA(cudaMemset(r->actions.p, 0, T * r->K * sizeof(int)));
    MZB_CUDA(r->d_slots.alloc(T * sizeof(int)));
    MZB_CUDA(r->d_push_actions.alloc(T * sizeof(int)));
    MZB_CUDA(r->d_push_frames.alloc(T * r->frame_elems * sizeof(float)));
    // config.action_representation_initial pre-evaluated per action: [A][ach * h * w]; NULL = the examples' constant plane a / A
    std::vector<float> planes((size_t)r->A * std::max(1, r->ach) * r->hw, 0.f);
    if (r->ach > 0) {
        if (action_planes_initial) std::copy(action_planes_initial, action_planes_initial + (size_t)r->A * r->ach * r->hw, planes.begin());
        else for (int a = 0; a < r->A; ++a) std::fill(planes.begin() + (size_t)a * r->ach * r->hw, planes.begin() + (size_t)(a + 1) * r->ach * r->hw, (float)((double)a / r->A));
    }
    MZB_CUDA(r->planes.alloc(planes.size() * sizeof(float)));
    MZB_CUDA(cudaMemcpy(r->planes.p, planes.data(), planes.size() * sizeof(float), cudaMemcpyHostToDevice));
    r->h_count.assign(T, 0);
    r->actor.resize(T);
    r->actor_seeded.assign(T, 0);
    e->ring = std::move(r);
    return MZB_OK;
}

int mzb_obs_ring_reset(mzb_engine* e, const int32_t* slots, int n) {
    int rc = ring_check(e, slots, n, false);
    if (rc) return rc;
    MZB_CUDA(cudaSetDevice(e->device));
    ObsRing& r = *e->ring;
    MZB_CUDA(cudaMemcpyAsync(r.d_slots.p, slots, (size_t)n * sizeof(int), cudaMemcpyHostToDevice, e->stream));
    ring_count_kernel<<<(n + 127) / 128, 128, 0, e->stream>>>(r.count.as<int>(), r.d_slots.as<int>(), n, 1);
    MZB_CUDA(cudaStreamSynchronize(e->stream));
    for (int i = 0; i < n; ++i) r.h_count[slots[i]] = 0;
    return MZB_OK;
}

int mzb_obs_ring_push(mzb_engine* e, const int32_t* slots, const float* frames, const int32_t* actions, int n) {
    int rc = ring_check(e, slots, n, false);
    if (rc) return rc;
    MZB_REQUIRE(frames && actions, "null argument");
    ObsRing& r = *e->ring;
    for (int i = 0; i < n; ++i) MZB_REQUIRE(r.h_count[slots[i]] == 0 || (actions[i] >= 0 && actions[i] < r.A), "row %d: action %d outside [0, %d)", i, actions[i], r.A);
    MZB_CUDA(cudaSetDevice(e->device));
    cudaStream_t s = e->stream;
    MZB_CUDA(cudaMemcpyAsync(r.d_slots.p, slots, (size_t)n * sizeof(int), cudaMemcpyHostToDevice, s));
    MZB_CUDA(cudaMemcpyAsync(r.d_push_actions.p, actions, (size_t)n * sizeof(int), cudaMemcpyHostToDevice, s));
    MZB_CUDA(cudaMemcpyAsync(r.d_push_frames.p, frames, (size_t)n * r.frame_elems * sizeof(float), cudaMemcpyHostToDevice, s));
    const int bx = (int)std::min<long>(32, (r.frame_elems + 255) / 256);
    ring_push_kernel<<<dim3(bx, n), 256, 0, s>>>(r.frames.as<float>(), r.actions.as<int>(), r.count.as<int>(), r.d_slots.as<int>(), r.d_push_frames.as<float>(),
                                                r.d_push_actions.as<int>(), r.K, r.frame_elems);
    ring_count_kernel<<<(n + 127) / 128, 128, 0, s>>>(r.count.as<int>(), r.d_slots.as<int>(), n, 0);
    MZB_CUDA(cudaGetLastError());
    MZB_CUDA(cudaStreamSynchronize(s));
    for (int i = 0; i < n; ++i) ++r.h_count[slots[i]];
    return MZB_OK;
}

int mzb_obs_ring_stacked(mzb_engine* e, const int32_t* slots, int n, float* out) {
    int rc = ring_check(e, slots, n, true);
    if (rc) return rc;
    MZB_REQUIRE(out, "null argument");
    MZB_CUDA(cudaSetDevice(e->device));
    rc = ring_stack_into(e, slots, n, e->d_obs.as<float>(), e->stream);
    if (rc) return rc;
    MZB_CUDA(cudaMemcpyAsync(out, e->d_obs.p, (size_t)n * e->dims.initial_flat() * sizeof(float), cudaMemcpyDeviceToHost, e->stream));
    MZB_CUDA(cudaStreamSynchronize(e->stream));
    return MZB_OK;
}

int mzb_search_batch_ring(mzb_engine* e, const int32_t* slots, const uint32_t* legal_mask, const int32_t* to_play,
                          const uint8_t* add_noise, int n, mzb_search_result* out) {
    int rc = ring_check(e, slots, n, true);
    if (rc) return rc;
    MZB_CUDA(cudaSetDevice(e->device));
    rc = ring_stack_into(e, slots, n, e->d_obs.as<float>(), e->stream);
    if (rc) return rc;
    return search_with_device_obs_slots(e, slots, legal_mask, to_play, add_noise, n, out);
}

int mzb_actor_seed(mzb_engine* e, const int32_t* slots, const int32_t* seeds, int n) {
    int rc = ring_check(e, slots, n, false);
    if (rc) return rc;
    MZB_REQUIRE(seeds, "null argument");
    for (int i = 0; i < n; ++i) { e->ring->actor[slots[i]].seed((uint32_t)seeds[i]); e->ring->actor_seeded[slots[i]] = 1; }
    return MZB_OK;
}

// select_opponent_action's Random branch (self_play.cpp:47-50): std::uniform_int_distribution<>(0, n_legal - 1) on the slot's actor
// generator — the same generator select_action draws from, as in the reference, so the streams stay in step
int mzb_actor_uniform_int(mzb_engine* e, const int32_t* slots, const int32_t* upper_inclusive, int n, int32_t* out) {
    int rc = ring_check(e, slots, n, false);
    if (rc) return rc;
    MZB_REQUIRE(upper_inclusive && out, "null argument");
    for (int i = 0; i < n; ++i) {
        MZB_REQUIRE(e->ring->actor_seeded[slots[i]], "slot %d: actor generator not seeded (mzb_actor_seed)", slots[i]);
        MZB_REQUIRE(upper_inclusive[i] >= 0, "row %d: empty range", i);
        std::uniform_int_distribution<> d(0, upper_inclusive[i]);
        out[i] = d(e->ring->actor[slots[i]]);
    }
    return MZB_OK;
}

// select_action (self_play.cpp:21-39), row i of a batch of search results, with slot i's actor generator
int mzb_select_actions(mzb_engine* e, const int32_t* slots, const mzb_search_result* stats, int n, const double* temperature, int32_t* actions_out) {
    int rc = ring_check(e, slots, n, false);
    if (rc) return rc;
    MZB_REQUIRE(stats && stats->n_children && stats->child_actions && stats->children_relative_visit && temperature && actions_out, "null argument");
    const int A = e->cfg.num_actions;
    for (int i = 0; i < n; ++i) {
        MZB_REQUIRE(e->ring->actor_seeded[slots[i]], "slot %d: actor generator not seeded (mzb_actor_seed)", slots[i]);
        const int L = stats->n_children[i];
        MZB_REQUIRE(L >= 1 && L <= A, "row %d: %d children", i, L);
        const double* rel = stats->children_relative_visit + (size_t)i * A;
        const int32_t* acts = stats->child_actions + (size_t)i * A;
        if (temperature[i] <= 0) {
            actions_out[i] = acts[std::distance(rel, std::max_element(rel, rel + L))];
        } else {
            std::vector<double> w(rel, rel + L);
            for (double& v : w) v = std::pow(v, 1.0 / temperature[i]);
            std::discrete_distribution<> d(w.begin(), w.end());
            actions_out[i] = acts[d(e->ring->actor[slots[i]])];
        }
    }
    return MZB_OK;
}

}  // extern "C"
