// engine.h — the engine object behind the C ABI (include/mzb200.h).
#pragma once
#include <map>
#include <memory>
#include <string>
#include <utility>
#include <vector>

#include "net_backend.h"
#include "tree.h"

namespace mzb {

// URBG view over a raw mt19937 state with the exact interface std::mt19937 shows to libstdc++'s distributions
// (result_type = uint_fast32_t, min 0, max 2^32-1), so std::gamma_distribution consumes it identically.
struct HostMt {
    typedef uint_fast32_t result_type;
    uint32_t* mt;
    int* idx;
    static constexpr result_type min() { return 0; }
    static constexpr result_type max() { return 0xffffffffUL; }
    result_type operator()() {
        if (*idx >= 624) {
            for (int k = 0; k < 624; ++k) {
                uint32_t y = (mt[k] & 0x80000000u) | (mt[(k + 1) % 624] & 0x7fffffffu);
                mt[k] = mt[(k + 397) % 624] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
            }
            *idx = 0;
        }
        uint32_t y = mt[(*idx)++];
        y ^= y >> 11; y ^= (y << 7) & 0x9d2c5680u; y ^= (y << 15) & 0xefc60000u; y ^= y >> 18;
        return y;
    }
};

}  // namespace mzb

using namespace mzb;

namespace mzb { struct ObsRing; void obs_ring_free(ObsRing*); }
struct ObsRingDeleter { void operator()(mzb::ObsRing* r) const { mzb::obs_ring_free(r); } };

inline bool is_synthetic(int evaluator) { return evaluator == MZB_EVAL_SYNTHETIC || evaluator == MZB_EVAL_SYNTHETIC_OBS; }

struct mzb_engine {
    mzb_config cfg;
    int device = 0;
    NetDims dims;
    EncoderConsts venc, renc;
    NetPlan plan;
    std::unique_ptr<NetBackend> net;
    bool weights_loaded = false;
    // what mzb_load_weights / mzb_load_checkpoint last received, un-folded (VPRNetModel::SaveCheckpoint writes these back out)
    struct RawTensor { std::string name; std::vector<int64_t> dims; std::vector<float> data; };
    std::vector<RawTensor> raw_weights;
    std::vector<std::pair<std::string, int64_t>> raw_longs;
    int64_t weights_serial = 0;                               // bumped by every load
    // training state of mzb_learn_forward (learn.cu): the module's tensors with BatchNorm running statistics as moved so far
    std::map<std::string, std::vector<float>> learn_state;
    int64_t learn_state_serial = -1;
    bool learn_host_dirty = false;                            // mzb_learn_forward moved statistics the device training state has not seen
    void* learn_step = nullptr;                               // device-resident training state of mzb_learn_step (learn_step.cu)
    void (*learn_step_free)(void*) = nullptr;
    std::vector<float> action_planes;   // owned copy of cfg.action_planes
    static constexpr int MAX_PARTS = 8;
    cudaStream_t stream = nullptr, aux = nullptr;
    cudaStream_t pipe[MAX_PARTS - 1] = {};   // streams of parts 1.. of the pipelined simulation loop (part 0 runs on `stream`)
    cudaEvent_t ev_start = nullptr, ev_root = nullptr, ev_end = nullptr, ev_aux = nullptr, ev_fork = nullptr;
    cudaEvent_t ev_join[MAX_PARTS - 1] = {};
    cudaEvent_t ev_tower[MAX_PARTS] = {};
    cudaStream_t hi[MAX_PARTS] = {};          // high-priority streams of the parts' tree steps (MZB_TREE_PRIORITY, created on first use)
    cudaEvent_t ev_tree[MAX_PARTS] = {};
    // CUDA graph of the pipelined simulation loop (engine_search.cu)
    cudaGraphExec_t sim_graph = nullptr; uint64_t sim_graph_key = 0, sim_graph_warm_key = 0; int64_t sim_graph_launches = 0; bool graph_failed = false;   // end of each part's latest tower kernel (alternating pipeline)
    // trees
    TreeDev td;
    DevBuf meta, prior, value_sum, reward, mm_min, mm_max, root_to_play, n_legal, n_slots, max_depth, mt, mt_idx,
        path_slot, path_len, in_slot, out_slot, action, value, reward_out, policy, pbc_table;
    DevBuf trace_parent, trace_action, rec_value, rec_reward, rec_policy;
    bool recording = false;
    DevBuf dbg_score;   // mzb_debug_select_child
    // search inputs / outputs on the device
    DevBuf d_obs, d_legal, d_to_play, d_add_noise, d_noise, d_root_slots, d_iota;
    DevBuf d_slots, mt_rows, mt_idx_rows;   // mzb_search_batch_slots: slot list and compact RNG rows
    DevBuf r_root_value, r_depth, r_nchild, r_actions, r_visits, r_rel, r_policy;
    // MZB_EVAL_HOST: caller-supplied Evaluator callbacks; host rows of the search in flight
    mzb_host_evaluator host_eval = {nullptr, nullptr, nullptr};
    const float* h_obs_rows = nullptr;
    const int32_t* h_slot_rows = nullptr;
    // SE stub
    SEDev se;
    DevBuf se_cid, se_pool, se_mv;   // se_mv: two (parent, child) slot-pair buffers for the state mover [2][2][T]
    cudaEvent_t ev_se_out[2] = {nullptr, nullptr}, ev_se_mv[2] = {nullptr, nullptr};
    // pinned host staging
    void* h_pin = nullptr;
    size_t h_pin_bytes = 0;
    // compat-path scratch
    DevBuf c_hidden_in, c_hidden_out, c_action, bench_slots;
    mzb_search_timing timing;
    // device-resident observation history + actor generators (selfplay.cu)
    std::unique_ptr<mzb::ObsRing, ObsRingDeleter> ring;
    // nccl (dlopen'ed)
    void* nccl_lib = nullptr;
    void* nccl_comm = nullptr;
    // two-phase weight hand-over (engine_nccl.cu): receive stream, shadow blob, 0 = nothing pending / 1 = sent (root) / 2 = received
    cudaStream_t sync_stream = nullptr;
    cudaEvent_t ev_sync = nullptr;
    DevBuf shadow_blob;
    int sync_pending = 0;
};


namespace mzb {
int learn_sync_host(mzb_engine* e);   // learn_step.cu: device training state -> e->learn_state (no-op when nothing is newer there)
int ensure_pinned(mzb_engine* e, size_t bytes);
}
