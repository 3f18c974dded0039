/* mzb200.h — C ABI of libmzb200.so, the B200-native replacement for the self-play search hot path of
 * tuero/muzero-cpp (lock-step batched MCTS + the initial/recurrent inference calls it drives).
 *
 * Plain C: pointers and sizes only, no torch / libtorch / C++ types.  Every entry point names the reference
 * interface it replaces (paths relative to the reference repository root).  All functions return 0 on success
 * and a negative mzb_status otherwise; mzb_last_error() gives the message (thread-local).  The reference has
 * no error codes of its own (assert / std::exit / c10::Error); the C++ facade turns a failure into an abort.
 *
 * Threading: an engine is bound to one CUDA device and is driven by one host thread at a time (the reference's
 * MCTS objects are thread-confined too, self_play.cpp:228-238); use one engine per GPU / per rank.
 */
#ifndef MZB200_H_
#define MZB200_H_
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MZB_MAX_ACTIONS 32
#define MZB_MAX_HEAD_LAYERS 8

typedef enum { MZB_OK = 0, MZB_ERR_INVALID = -1, MZB_ERR_CUDA = -2, MZB_ERR_STATE = -3, MZB_ERR_NCCL = -4 } mzb_status;

/* arithmetic mode of the network towers */
typedef enum {
    MZB_FP32_CHECK = 0, /* plain fp32 CUDA-core kernels: <= 1e-4 abs vs libtorch fp32 (north star) */
    MZB_BF16 = 1        /* tcgen05 implicit-GEMM towers, bf16 operands / fp32 accumulate: <= 2e-2 rel */
} mzb_precision;

/* what stands behind Evaluator::{Initial,Recurrent}Inference (muzero-cpp/vprnet_evaluator.h:35-41) */
typedef enum {
    MZB_EVAL_NETWORK = 0,   /* the VPRNet towers (vprnet.cpp / models.cpp / model_layers.cpp) */
    MZB_EVAL_SYNTHETIC = 1, /* deterministic hash evaluator "SE": pins the tree kernels bit-exactly and measures them
                               alone (the stub still moves one hidden state parent->child per simulation) */
    MZB_EVAL_HOST = 2,      /* a caller-supplied Evaluator subclass (e.g. tests/mcts_test.cpp:16-27 TestEvaluator): the tree
                               kernels run on the GPU, the two inference calls go to host callbacks once per cohort and
                               simulation (mzb_set_host_evaluator).  Compatibility / test path, not a throughput path. */
    MZB_EVAL_SYNTHETIC_OBS = 3 /* "SE" keyed by a hash of the WHOLE stacked observation (rows of initial_flat_size floats, like
                               the network's): lets real games — play_game / reanalyze over AbstractGame, the observation ring —
                               be compared decision for decision with the reference driven by the same evaluator */
} mzb_evaluator;

/* POD mirror of muzero_config::MuZeroConfig + MuZeroNetworkConfig (muzero-cpp/config.h:15-117), restricted to the
 * fields the hot path reads.  Field names follow the reference. */
typedef struct {
    /* Game (config.h:52-66) */
    int32_t observation_c, observation_h, observation_w; /* observation_shape */
    int32_t num_actions;                                 /* action_space = {0..num_actions-1} */
    int32_t action_channels;
    int32_t num_players;                                 /* 1 or 2 (mcts.cpp:30) */
    int32_t stacked_observations;
    double value_upperbound, value_lowerbound;           /* "unknown" = (lowest, max) sentinels, types.h:77-78 */
    /* categorical transform (config.h:69-73) */
    double min_reward, max_reward, min_value, max_value;
    int32_t use_contractive_mapping;
    /* search (config.h:83-93) */
    int32_t num_simulations;
    double discount, dirichlet_alpha, dirichlet_epsilon, pb_c_base, pb_c_init;
    int32_t seed;                                        /* tree slot i is seeded seed + i (self_play.cpp:232) */
    /* network_config (config.h:15-31) */
    int32_t downsample, normalize_hidden_states;
    int32_t resnet_channels, representation_blocks, dynamics_blocks, prediction_blocks;
    int32_t reward_reduced_channels, policy_reduced_channels, value_reduced_channels;
    int32_t n_reward_head_layers, n_policy_head_layers, n_value_head_layers;
    int32_t reward_head_layers[MZB_MAX_HEAD_LAYERS], policy_head_layers[MZB_MAX_HEAD_LAYERS],
        value_head_layers[MZB_MAX_HEAD_LAYERS];
    /* action_representation_recurrent (types.h:43, vprnet.cpp:143) pre-evaluated once per action:
     * [num_actions][action_channels * enc_h * enc_w] floats; NULL = the examples' constant plane action/num_actions */
    const float* action_planes;
    /* engine */
    int32_t max_trees;    /* concurrent tree slots resident on this GPU */
    int32_t precision;    /* mzb_precision */
    int32_t evaluator;    /* mzb_evaluator */
    double se_value_lo, se_value_hi, se_reward_lo, se_reward_hi; /* SE output ranges (evaluator == SYNTHETIC) */
} mzb_config;

typedef struct mzb_engine mzb_engine;

/* One named fp32 tensor of the reference module archive (checkpoint-<step>.pt), e.g.
 * "dynamics_network.dynamics_resnet_body.0.resnet_0_conv1.weight", "...resnet_0_bn1.running_var". */
typedef struct {
    const char* name;
    const float* data;
    int32_t ndim;
    int64_t dims[4];
} mzb_tensor_desc;

/* SoA result of a batch of searches = T x types::MCTSReturn (muzero-cpp/types.h:70-76).  Caller-owned host arrays;
 * any pointer may be NULL to skip that output.  Rows are padded to num_actions. */
typedef struct {
    double* root_value;               /* [T] */
    int32_t* max_tree_depth;          /* [T] */
    int32_t* n_children;              /* [T] number of legal root actions L */
    int32_t* child_actions;           /* [T][A] first L valid, in the FINAL SHUFFLED child order (mcts.cpp:187) */
    int32_t* child_visits;            /* [T][A] integer visit counts, same order */
    double* children_relative_visit;  /* [T][A] visits / sum (mcts.cpp:192-194) */
    double* emperical_policy;         /* [T][A] dense, indexed by action id (mcts.cpp:196-199) */
} mzb_search_result;

const char* mzb_last_error(void);
const char* mzb_version(void);

/* Replaces: VPRNetModel::VPRNetModel (vprnet.cpp:30-50) + DeviceManager::AddDevice (device_manager.h:36-39) +
 * MCTS::MCTS (mcts.cpp:17-31) for max_trees tree slots.  device = CUDA ordinal. */
int mzb_engine_create(const mzb_config* config, int device, mzb_engine** out);
void mzb_engine_destroy(mzb_engine* e);

/* MZB_EVAL_HOST: the two pure virtuals of Evaluator (vprnet_evaluator.h:35-41), batched over the rows of a cohort.
 * slot[i] = tree slot of row i.  Nodes are named by expansion order (root 0; the node expanded by simulation s is s + 1), so
 * the callee can keep each node's encoded_state itself: `recurrent` expands node `child_node` of every row under
 * parent_node[i] by action[i].  Outputs: value[n], reward[n], policy[n][num_actions].  Non-zero return aborts the search. */
typedef struct {
    void* user;
    int (*initial)(void* user, int n, const int32_t* slot, const float* obs, int obs_stride, double* value, double* reward,
                   double* policy);
    int (*recurrent)(void* user, int n, const int32_t* slot, const int32_t* parent_node, const int32_t* action, int child_node,
                     double* value, double* reward, double* policy);
} mzb_host_evaluator;
int mzb_set_host_evaluator(mzb_engine* e, const mzb_host_evaluator* evaluator);

/* Replaces: VPRNetModel::LoadCheckpoint (vprnet.cpp:66-74): takes the module's named fp32 tensors (78 params + 60
 * buffers for the shipped Connect4 net), folds eval-mode BN (eps 1e-4) and packs them for the kernels. */
int mzb_load_weights(mzb_engine* e, const mzb_tensor_desc* tensors, int n_tensors);

/* Replaces: the read half of VPRNetModel::LoadCheckpoint(path) (vprnet.cpp:71-74): reads the module archive
 * torch::save(model_, "<path>.pt") wrote (vprnet.cpp:58-63; a ZIP of stored members with a protocol-2 pickle of the module tree)
 * natively — no libtorch — and loads its float32 parameters and buffers like mzb_load_weights.  `path` is the file itself
 * (the facade appends ".pt" like the reference).  The "-optimizer.pt" file belongs to the learner and is not read.
 * mzb_checkpoint_open/count/tensor/close expose the parsed tensors (names as in the archive, row-major float32 copies valid
 * until close) without touching a GPU. */
typedef struct mzb_checkpoint mzb_checkpoint;
int mzb_checkpoint_open(const char* path, mzb_checkpoint** out);
int mzb_checkpoint_count(const mzb_checkpoint* checkpoint);
int mzb_checkpoint_tensor(const mzb_checkpoint* checkpoint, int index, mzb_tensor_desc* out);
void mzb_checkpoint_close(mzb_checkpoint* checkpoint);
int mzb_load_checkpoint(mzb_engine* e, const char* path);
/* int64 scalars of the archive (BatchNorm num_batches_tracked; eval-mode inference never reads them, a faithful re-save does) */
int mzb_checkpoint_long_count(const mzb_checkpoint* checkpoint);
int mzb_checkpoint_long(const mzb_checkpoint* checkpoint, int index, const char** name, int64_t* value);

/* Replaces: the write half, VPRNetModel::SaveCheckpoint(step) = torch::save(model_, "<path>.pt") (vprnet.cpp:58-63; SURVEY §8f
 * rank 3).  Emits the same module archive natively — ZIP of stored members, protocol-2 pickle of the module tree rebuilt from the
 * dotted tensor names, one TorchScript class definition per module — so that a reference process (torch::load, vprnet.cpp:71-74)
 * or `--resume` reads what this engine holds.  mzb_checkpoint_write: any set of named float32 tensors (+ optional int64 scalars
 * "....num_batches_tracked"; BatchNorm modules without one get 0), no GPU involved.  mzb_save_checkpoint: the tensors the engine
 * was last loaded with (mzb_load_weights / mzb_load_checkpoint; NOT what a later mzb_broadcast_weights received into the packed
 * blob).  The file appears atomically (written beside `path`, then renamed).  The optimiser file is the learner's. */
int mzb_checkpoint_write(const char* path, const mzb_tensor_desc* tensors, int n_tensors, const char* const* long_names,
                         const int64_t* long_values, int n_longs);
int mzb_save_checkpoint(mzb_engine* e, const char* path);

/* Replaces (first slice, forward half): VPRNetModel::Learn (vprnet.cpp:192-311) up to the loss, SURVEY.md 8f rank 2.
 * mzb_batch = types::Batch (types.h:81-90): B samples, each with K + 1 = num_unroll_steps + 1 steps.  mzb_learn_forward runs the
 * K + 1 unrolled network calls in TRAIN mode (batch-statistics BatchNorm, eps 1e-4; the running statistics move by momentum 0.01 and
 * persist across calls) on the tensors last loaded into the engine, encodes the value / reward targets (ValueEncoder::encode,
 * util.cpp:59-76) and evaluates MuzeroNetworkImpl::loss (models.cpp:268-276) — fp32 on the GPU, whatever the engine's precision mode.
 * mzb_loss_info = VPRNetModel::LossInfo (vprnet.h:33-50): total_loss = mean(priorities [B,1] * per-sample loss [B]) as the reference
 * broadcasts it, value / policy / reward = batch means of the unscaled terms, errors[B] = decoded step-0 value - target_values[:,0]
 * (may be NULL).  No backward pass and no optimiser step yet: weights are not changed.  gradient_scale is carried for that later
 * half.  mzb_learn_get_tensor reads a tensor of the training state (e.g. a BatchNorm running_var after the call). */
typedef struct {
    int32_t num_samples;                 /* B */
    const int32_t* actions;              /* [B][K+1] */
    const float* stacked_observations;   /* [B][initial_flat_size] */
    const double* target_values;         /* [B][K+1] */
    const double* target_rewards;        /* [B][K+1] */
    const double* target_policies;       /* [B][K+1][A] */
    const double* gradient_scale;        /* [B]; backward only, may be NULL */
    const double* priorities;            /* [B]: importance-sampling weights */
} mzb_batch;
typedef struct {
    double total_loss, value, policy, reward;
    double* errors;                      /* [B] or NULL */
} mzb_loss_info;
int mzb_learn_forward(mzb_engine* e, const mzb_batch* batch, int num_unroll_steps, double value_loss_weight, mzb_loss_info* out);
int mzb_learn_get_tensor(mzb_engine* e, const char* name, float* out, int64_t numel);
/* Replaces: the whole of VPRNetModel::Learn (vprnet.cpp:192-311) — forward as mzb_learn_forward, then total_loss.backward() with
 * Learn's gradient hooks (x0.5 on every recurrent hidden state, gradient_scale on the losses of unroll steps >= 1) and
 * model_optimizer_.step() = torch::optim::Adam(lr, weight_decay = l2) (vprnet.cpp:35-36; betas 0.9 / 0.999, eps 1e-8), fp32 on the GPU.
 * Parameters, gradients and Adam moments stay resident on the device between calls; a load (mzb_load_weights / _checkpoint) starts the
 * optimizer over, like constructing a new VPRNetModel.  mzb_learn_get_tensor reads a trained tensor back, mzb_learn_get_gradient
 * the gradient of the last step.  mzb_learn_commit = the learner -> actor hand-over on one GPU (learner.cpp:93-101): the trained
 * tensors become the weights the engine searches with and mzb_save_checkpoint writes. */
int mzb_learn_step(mzb_engine* e, const mzb_batch* batch, int num_unroll_steps, double value_loss_weight, double learning_rate,
                   double l2_weight_decay, mzb_loss_info* out);
int mzb_learn_get_gradient(mzb_engine* e, const char* name, float* out, int64_t numel);
int mzb_learn_commit(mzb_engine* e);

/* Replaces: PrioritizedReplayBuffer (replay_buffer.cpp:15-141) over SumTree (replay_buffer.h:28-209) and the target construction of
 * GameHistory (compute_target_value / make_target / get_stacked_observations, types.cpp:16-95) — SURVEY.md 8f rank 4.  The sum tree
 * and the leaves' (step, history) stay on the host; the histories themselves (frames, actions, rewards, to_play, root values, visit
 * distributions) are stored in HBM when a game is saved, and one kernel launch assembles a whole training batch from them.  Random
 * draws come from the CALLER's generator (`next` returns its 32-bit outputs, e.g. std::mt19937::operator()) through libstdc++'s own
 * distributions in the reference's order, so that the same generator state yields the same batch as the reference, bit for bit.
 * mzb_game_history: length L = root_values.size(); observations [L+1][c*h*w], actions / rewards / to_play [L+1] (action_history[0]
 * = 0 as play_game pushes it), root_values [L], child_visits [L][A].  mzb_batch_out = types::Batch with caller-allocated arrays:
 * priorities (importance weights) / indices (tree leaves) / gradient_scale [B], actions / target_values / target_rewards [B][K+1],
 * target_policies [B][K+1][A], stacked_observations [B][initial_flat_size]. */
typedef uint32_t (*mzb_rng_fn)(void* user);
typedef struct {
    int32_t capacity;                    /* replay_buffer_size: stored positions */
    int32_t min_sample_size, td_steps, num_unroll_steps;
    double per_alpha, per_beta, per_epsilon, per_beta_increment;
    const float* action_planes_initial;  /* action_representation_initial per action [A][ach*h*w]; NULL = constant plane a / A */
} mzb_replay_config;
typedef struct {
    int32_t length;
    const float* observations;
    const int32_t* actions;
    const double* rewards;
    const int32_t* to_play;
    const double* root_values;
    const double* child_visits;
} mzb_game_history;
typedef struct {
    int32_t num_samples;
    double* priorities; int32_t* indices; int32_t* actions; float* stacked_observations;
    double* target_rewards; double* target_values; double* target_policies; double* gradient_scale;
} mzb_batch_out;
typedef struct mzb_replay mzb_replay;
int mzb_replay_create(mzb_engine* e, const mzb_replay_config* config, mzb_replay** out);
void mzb_replay_destroy(mzb_replay* r);
int mzb_replay_save_game_history(mzb_replay* r, const mzb_game_history* history);
int mzb_replay_size(const mzb_replay* r);
int mzb_replay_can_sample(const mzb_replay* r);
int mzb_replay_num_histories(const mzb_replay* r);
double mzb_replay_total_priority(const mzb_replay* r);
int mzb_replay_oldest_history(const mzb_replay* r);   /* ids below it have been evicted (first in, first out) */
int mzb_replay_sample(mzb_replay* r, mzb_rng_fn next, void* user, int batch_size, mzb_batch_out* out);
int mzb_replay_update_priorities(mzb_replay* r, const int32_t* indices, const double* errors, int n);
int mzb_replay_sample_game(mzb_replay* r, mzb_rng_fn next, void* user, int32_t* history_id, int32_t* length);
int mzb_replay_update_game_history(mzb_replay* r, int history_id, const double* root_values, const double* child_visits, int length);

/* Replaces: the disk round trip SaveCheckpoint -> LoadCheckpoint of learner.cpp:93-101 / muzero.cpp:142-148.
 * One ncclBroadcast of the packed weight blob from `root` to all ranks.  id = 128-byte ncclUniqueId. */
int mzb_nccl_unique_id(char id_out[128]);
int mzb_nccl_init(mzb_engine* e, const char id[128], int rank, int world_size);
int mzb_broadcast_weights(mzb_engine* e, int root);
/* The same hand-over in two phases, so that a learner's broadcast overlaps self-play.  _begin: the blob travels on a
 * stream of its own into a SHADOW buffer (the root only sends; `rank` = this engine's rank in the communicator) while searches keep
 * running on the live weights; returns without waiting.  mzb_weights_commit: enqueued on the search stream — waits for the receive
 * and installs the shadow copy between two searches (stream order), never under one.  Commit without a pending _begin is a no-op. */
int mzb_broadcast_weights_begin(mzb_engine* e, int root, int rank);
int mzb_weights_commit(mzb_engine* e);
/* checksum (FNV-1a over the packed blob, read back from the device) to verify a broadcast */
int mzb_weights_checksum(mzb_engine* e, uint64_t* out);

/* Replaces: VPRNetModel::InitialInference (vprnet.cpp:77-129).  Host pointers, exactly the reference's batch API:
 * obs [B][Cin*H*W] -> value[B], reward[B] (== 0), policy[B][A] (softmax), hidden[B][C*h*w] (NCHW order, fp32). */
int mzb_initial_inference(mzb_engine* e, const float* obs, int batch, double* value, double* reward, double* policy,
                          float* hidden);
/* Replaces: VPRNetModel::RecurrentInference (vprnet.cpp:131-189).  hidden [B][C*h*w] NCHW, action [B]. */
int mzb_recurrent_inference(mzb_engine* e, const float* hidden, const int32_t* action, int batch, double* value,
                            double* reward, double* policy, float* hidden_out);
/* recurrent_inference states/s with device-resident inputs (BASELINE metric M2): runs `iters` passes over an
 * internal batch of `batch` random hidden states and reports the mean device time per pass (CUDA events). */
int mzb_bench_recurrent(mzb_engine* e, int batch, int iters, float* ms_per_pass, int* launches_per_pass);

/* Replaces: MCTS::MCTS's rng_(seed) (mcts.cpp:26) for tree slots [first, first+n). */
int mzb_search_seed(mzb_engine* e, int first_slot, int n, const int32_t* seeds);

/* Replaces: T concurrent MCTS::run calls (mcts.cpp:34-87; callers self_play.cpp:130,208), tree slot t <- row t.
 * obs [T][Cin*H*W] host (SYNTHETIC evaluator: obs row = one float, the position key), legal_mask[T] bit a = action a
 * legal, to_play[T], add_noise[T].  Blocks until all T searches finished.
 * The root's children are created in ASCENDING action order.  The reference creates them in the order of the caller's
 * legal_actions vector (mcts.cpp:163-166), and that order feeds std::shuffle and the first-strict-maximum tie rule: the two agree
 * bit for bit when the vector is strictly ascending and duplicate-free — what every AbstractGame::legal_actions() of the reference's
 * examples returns; the C++ headers above this ABI (legal_mask_of) refuse any other vector. */
int mzb_search_batch(mzb_engine* e, const float* obs, const uint32_t* legal_mask, const int32_t* to_play,
                     const uint8_t* add_noise, int n_trees, mzb_search_result* out);
/* Same search with observations already resident in HBM (device pointer, engine layout = row-major fp32). */
int mzb_search_batch_device_obs(mzb_engine* e, const float* d_obs, const uint32_t* legal_mask, const int32_t* to_play,
                                const uint8_t* add_noise, int n_trees, mzb_search_result* out);
/* Same search for an arbitrary subset of tree slots: row i uses (and advances) the RNG stream of slot slots[i]
 * (distinct, in [0, max_trees)).  This is what the blocking one-tree MCTS::run facade (include/mzb200_facade.hpp)
 * batches through: actors whose games finished simply stop submitting their slot. */
int mzb_search_batch_slots(mzb_engine* e, const float* obs, const int32_t* slots, const uint32_t* legal_mask,
                           const int32_t* to_play, const uint8_t* add_noise, int n_trees, mzb_search_result* out);
/* device time (CUDA events) of the last search, split as far as the engine can tell */
typedef struct {
    float total_ms, root_ms, sim_loop_ms;
    int64_t kernel_launches;  /* kernels launched (graph nodes included) */
    int64_t simulations;      /* T * num_simulations */
} mzb_search_timing;
int mzb_last_search_timing(mzb_engine* e, mzb_search_timing* out);

/* Parity hooks.  Enable before a search; afterwards read, per tree, the expansion trace ("tree indices": node i+1
 * was expanded by simulation i under (parent node, action)) and the network outputs each node was expanded with
 * (node 0 = root) so the oracle can replay "identical network outputs". */
int mzb_search_set_recording(mzb_engine* e, int enabled);
int mzb_search_get_trace(mzb_engine* e, int n_trees, int32_t* parent /*[T][n]*/, int32_t* action /*[T][n]*/);
int mzb_search_get_records(mzb_engine* e, int n_trees, double* value /*[T][n+1]*/, double* reward /*[T][n+1]*/,
                           double* policy /*[T][n+1][A]*/);

/* Unit-level parity hooks for the reference's hand-built trees (tests/mcts_test.cpp:30-153 drives MCTS::backpropogate,
 * select_child and ucb_score on MCTSNode objects it assembles itself).  A tree is uploaded into tree slot `slot` in the engine's
 * own layout — nodes in slot order, the children of a node contiguous, root = node 0 — one device step runs on it, and it is
 * read back.  meta rows are {visit_count, action, child_base (-1 = not expanded), tag}; the tag travels with its record through
 * the physical child shuffle of select_child (mcts.cpp:112).  Nodes below the root are taken to have num_actions children, as
 * MCTS::run creates them (mcts.cpp:72). */
typedef struct {
    int32_t n_nodes;
    const int32_t* meta;                        /* [n_nodes][4] */
    const double *prior, *value_sum, *reward;   /* [n_nodes] */
    double minmax_minimum, minmax_maximum;      /* MinMaxStats state (mcts.h:79-112) */
    int32_t root_to_play, n_root_children;
} mzb_debug_tree;
int mzb_debug_tree_write(mzb_engine* e, int slot, const mzb_debug_tree* tree);
int mzb_debug_tree_read(mzb_engine* e, int slot, int n_nodes, int32_t* meta, double* prior, double* value_sum, double* reward,
                        double* minmax /*[2] = minimum, maximum*/);
/* MCTS::select_child(root) (mcts.cpp:110-124) on the uploaded tree with the slot's RNG stream: shuffles the root's child
 * records in place and returns the node index of the chosen child; scores[j] = ucb_score(root, child j) (mcts.cpp:127-141) in
 * the post-shuffle order (num_actions entries, first n_root_children valid).  restore_rng != 0 puts the slot's generator back
 * afterwards (ucb_score itself draws nothing). */
int mzb_debug_select_child(mzb_engine* e, int slot, int restore_rng, int32_t* chosen_node, double* scores);
/* MCTS::backpropogate(search_path, value, to_play) (mcts.cpp:90-107) along path[0..len) (node indices, root first; the leaf's
 * to_play is root_to_play + len - 1 mod players); the device step also expands the leaf as MCTS::run would (reward
 * leaf_reward, num_actions children with `policy` priors appended at node index n_nodes).  The MinMaxStats state is updated. */
int mzb_debug_backpropagate(mzb_engine* e, int slot, const int32_t* path, int len, double value, double leaf_reward,
                            const double* policy);

/* MZB_EVAL_SYNTHETIC only: the hidden state the stub holds for node `node` (expansion order, 0 = root) of tree row `tree` after a
 * search — recurrent_flat_size floats.  The stub moves one state parent -> child per simulation (se_move_kernel), so every
 * expanded node must hold a copy of its tree's root state; tests check the mover through this. */
int mzb_debug_se_state(mzb_engine* e, int tree, int node, float* out);

/* ---- The callers' side: observation history and action selection for cohorts of games ----
 * Device-resident observation history per tree slot, stacked on the GPU the way GameHistory::get_stacked_observations
 * (types.cpp:16-42) does on the host: current frame, then for each of the previous `stacked_observations` steps the frame and
 * the planes of the action taken from it, zero planes where the game is younger.  play_game (self_play.cpp:84-160) maps to:
 *   reset -> mzb_obs_ring_reset;  game_history.observation_history/action_history.push_back -> mzb_obs_ring_push (one new frame
 *   per game and move crosses PCIe, not the stacked observation);  get_stacked_observations(-1, ...) + mcts.run ->
 *   mzb_search_batch_ring;  select_action -> mzb_select_actions.
 * action_planes_initial: config.action_representation_initial pre-evaluated, [num_actions][action_channels * h * w] at
 * OBSERVATION resolution; NULL = the shipped examples' constant plane action / num_actions. */
int mzb_obs_ring_create(mzb_engine* e, const float* action_planes_initial);
int mzb_obs_ring_reset(mzb_engine* e, const int32_t* slots, int n);
/* frames [n][c*h*w] (one raw observation per game); actions[i] = the action that led to it (ignored for a game's first frame:
 * the reference stores a dummy 0 there, self_play.cpp:91) */
int mzb_obs_ring_push(mzb_engine* e, const int32_t* slots, const float* frames, const int32_t* actions, int n);
/* the stacked observations of the slots' current positions, copied to the host (checking / callers that want the tensor) */
int mzb_obs_ring_stacked(mzb_engine* e, const int32_t* slots, int n, float* out /*[n][initial_flat_size]*/);
/* mzb_search_batch_slots with the observations taken from the ring (stacked on the device, no host copy) */
int mzb_search_batch_ring(mzb_engine* e, const int32_t* slots, const uint32_t* legal_mask, const int32_t* to_play,
                          const uint8_t* add_noise, int n_trees, mzb_search_result* out);
/* select_action (self_play.cpp:21-39): row i of `stats` with slot i's actor generator — std::mt19937(config.seed + actor_num),
 * self_play.cpp:233, a stream of its own next to the search's.  temperature <= 0: first maximum of the relative visits;
 * otherwise one std::discrete_distribution draw over visits^(1/temperature). */
int mzb_actor_seed(mzb_engine* e, const int32_t* slots, const int32_t* seeds, int n);
int mzb_select_actions(mzb_engine* e, const int32_t* slots, const mzb_search_result* stats, int n, const double* temperature,
                       int32_t* actions_out);
/* select_opponent_action's Random branch (self_play.cpp:47-50): out[i] = std::uniform_int_distribution<>(0, upper_inclusive[i]) drawn
 * from slot i's actor generator (the one mzb_select_actions uses, as the reference shares `rng` between the two). */
int mzb_actor_uniform_int(mzb_engine* e, const int32_t* slots, const int32_t* upper_inclusive, int n, int32_t* out);

/* ValueEncoder::encode (util.cpp:59-76) and ::decode (util.cpp:80-87) of the engine's value (which = 0) or reward (which = 1)
 * encoder, on the GPU, host pointers in and out: n scalars <-> n rows of support_size floats (mzb_shapes gives the sizes).
 * decode is what the head kernels fuse during inference; encode builds categorical targets (VPRNetModel::Learn's callers,
 * vprnet.cpp:240-243) and backs the reference's own round-trip tests (tests/util_test.cpp). */
int mzb_value_encode(mzb_engine* e, int which, const float* x, int n, float* support /*[n][S]*/);
int mzb_value_decode(mzb_engine* e, int which, const float* logits /*[n][S]*/, int n, int take_softmax, float* x /*[n]*/);
/* No counterpart in the reference: added for callers that cannot call cudaMalloc (CUDA graph
 * capture, tight serving loops, harnesses that forbid it).  mzb_value_decode with the same arguments, the same kernel and the same
 * results, but it allocates nothing: its device staging (n rows of S logits + n decoded scalars) lives in a file-scope
 * `static __device__` array of MZB_VALUE_DECODE_STATIC_MAX_FLOATS floats (4 MiB, held by every GPU that loads the library until the
 * process exits).  n * (S + 1) beyond that is rejected with MZB_ERR_INVALID and a message naming both numbers.  The array is shared by
 * every call on a device and keeps the last call's contents (nothing in it needs clearing: each call overwrites what it reads);
 * calls that overlap on the same GPU race on it — serialise them. */
#define MZB_VALUE_DECODE_STATIC_MAX_FLOATS (1 << 20)
int mzb_value_decode_static(mzb_engine* e, int which, const float* logits /*[n][S]*/, int n, int take_softmax, float* x /*[n]*/);

/* derived shapes (VPRNetModel ctor, vprnet.cpp:41-46; ValueEncoder::get_support_size, util.cpp:91-102) */
typedef struct {
    int32_t initial_channels, initial_flat_size; /* Cin, Cin*H*W */
    int32_t enc_c, enc_h, enc_w, recurrent_flat_size;
    int32_t value_support_size, reward_support_size;
    double value_step, value_min, reward_step, reward_min; /* ValueEncoder step_size_/min_value_ after h() */
    int64_t weight_blob_bytes;
    int64_t node_pool_bytes;
} mzb_shapes;
int mzb_get_shapes(mzb_engine* e, mzb_shapes* out);

#ifdef __cplusplus
}
#endif
#endif /* MZB200_H_ */
