// tree.h — structure-of-arrays search trees resident in HBM and the warp-per-tree kernels that walk them.
// Replaces muzero-cpp/mcts.{h,cpp}: MCTSNode (mcts.h:19-75) becomes parallel arrays indexed by [tree][slot];
// MCTS::run's simulation loop (mcts.cpp:52-79) becomes one tree_step launch per simulation, lock-step over all
// trees, between two network launches.  All f64 arithmetic keeps the reference's operation order and is compiled
// without FMA contraction so decisions are bit-identical to the x86-64 reference.
#pragma once
#include <cstdint>

#include <cuda_runtime.h>

namespace mzb {

struct TreeDev {
    int T, A, n_sims, slots_per_tree, players;
    // per slot, [T][slots_per_tree]; the children of a node are contiguous slots
    int4* meta;          // {visit_count, action, child_base (-1 = not expanded), node_id (-1 = not expanded)}
    double* prior;
    double* value_sum;
    double* reward;
    // per tree
    double* mm_min; double* mm_max;     // MinMaxStats (mcts.h:79-112)
    int* root_to_play; int* n_legal; int* n_slots; int* max_depth;
    uint32_t* mt; int* mt_idx;          // std::mt19937 state [T][624] + position
    int* path_slot;                     // [T][n_sims + 2]
    int* path_len;                      // [T]
    // network I/O for the current simulation
    int* in_slot; int* out_slot; int* action;      // [T] pool slots (parent state, child state) and action
    double* value; double* reward_out; double* policy;   // [T], [T], [T][A]
    // constants
    const double* pbc_table;            // [2][n_sims + 2]: log((N + pb_c_base + 1) / pb_c_base) + pb_c_init, then sqrt(N); host (glibc) evaluated
    double discount, init_min, init_max, dirichlet_epsilon;
    // optional recording (parity hooks); null when disabled
    int* trace_parent; int* trace_action;                  // [T][n_sims]
    double* rec_value; double* rec_reward; double* rec_policy;   // [T][n_sims+1], [T][n_sims+1], [T][n_sims+1][A]
    double* dbg_score;                                     // [32]: pUCT scores of the root's children (mzb_debug_select_child only)
};

// root set-up after the initial inference: expand(legal), optional noise, minmax reset (mcts.cpp:39-49)
void launch_tree_root(const TreeDev& td, int n_trees, const uint32_t* d_legal_mask, const int* d_to_play,
                      const uint8_t* d_add_noise, const double* d_noise /*[T][A]*/, cudaStream_t s);
// simulation step `sim` in [0, n_sims]:  (sim > 0: expand + backup of the previous selection)  then
// (sim < n_sims: select a leaf and publish the network request)
void launch_tree_step(const TreeDev& td, int n_trees, int sim, cudaStream_t s);
// the same step for trees [t_first, t_end) only: the two half-cohorts of the pipelined simulation loop
void launch_tree_step_range(const TreeDev& td, int t_first, int t_end, int sim, cudaStream_t s);
// get_search_statistics (mcts.cpp:179-200) into device SoA result arrays
struct ResultDev {
    double* root_value; int* max_tree_depth; int* n_children; int* child_actions; int* child_visits;
    double* relative_visit; double* emperical_policy;
};
void launch_tree_stats(const TreeDev& td, int n_trees, const ResultDev& r, cudaStream_t s);

// SE synthetic evaluator (MZB_EVAL_SYNTHETIC): stands in for the two network calls.
struct SEDev {
    uint32_t* node_cid;      // [T][n_sims + 1]
    float* pool;             // [T * (n_sims + 1)][slot_floats]: one hidden state is moved parent -> child per simulation
    int slot_floats;
    double vlo, vhi, rlo, rhi;
    int bulk;                // 1 = the state moves by bulk copy (se_move_kernel), 0 = per-lane copy inside the outputs kernel
    int* mv_in; int* mv_out; // [T]: this simulation's (parent slot, child slot) pairs, snapshotted for the state mover, which runs on
                             // its own stream beside the next tree step (two such buffers alternate)
};
void launch_se_initial(const TreeDev& td, const SEDev& se, int n_trees, const float* d_obs, int obs_stride, cudaStream_t s);
// stub outputs (+ the per-lane state copy for state sizes the bulk copy cannot take); snapshots the slot pairs into se.mv_in / mv_out
void launch_se_recurrent(const TreeDev& td, const SEDev& se, int n_trees, cudaStream_t s);
// the bulk-copy state mover for the pairs in se.mv_in / mv_out; false when the state size needs the per-lane copy instead
bool se_move_available(int slot_floats);
void launch_se_move(const SEDev& se, int n_trees, cudaStream_t s);

// RNG streams of a subset of slots <-> compact rows (mzb_search_batch_slots): dir 0 = gather rows <- slots, 1 = scatter back
void launch_mt_move(uint32_t* d_mt_home, int* d_idx_home, uint32_t* d_mt_rows, int* d_idx_rows, const int* d_slots, int n,
                    int dir, cudaStream_t s);

// seeds -> mt19937 states (std::mt19937 seeding recurrence), on the device
void launch_mt_seed(uint32_t* d_mt, int* d_idx, const int* d_seeds, int first_slot, int n, cudaStream_t s);

}  // namespace mzb
