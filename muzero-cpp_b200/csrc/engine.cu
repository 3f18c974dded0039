// engine.cu — the engine object behind the C ABI (include/mzb200.h): owns the trees, the node pool, the network
// backend and the lock-step search loop.  Host C++ only drives launches; no libtorch anywhere.
//
// Replaces, for its GPU: MCTS objects (mcts.cpp:17-31) for max_trees slots, VPRNetEvaluator's batching queues
// (vprnet_evaluator.cpp:105-178: lock-step cohorts need no queue, promise or 1 ms deadline), DeviceManager's
// model loan (device_manager.h:68-80: one resident model per engine), VPRNetModel's inference entry points.
#include <dlfcn.h>

#include <chrono>
#include <cmath>
#include <cstring>
#include <memory>
#include <numeric>
#include <random>

#include "engine.h"

namespace mzb {
int ensure_pinned(mzb_engine* e, size_t bytes) {
    if (bytes <= e->h_pin_bytes) return MZB_OK;
    if (e->h_pin) cudaFreeHost(e->h_pin);
    e->h_pin = nullptr;
    e->h_pin_bytes = 0;
    MZB_CUDA(cudaMallocHost(&e->h_pin, bytes));
    e->h_pin_bytes = bytes;
    return MZB_OK;
}
}  // namespace mzb

using namespace mzb;

namespace {

int alloc_trees(mzb_engine* e) {
    const mzb_config& c = e->cfg;
    const int T = c.max_trees, A = c.num_actions, n = c.num_simulations;
    TreeDev& td = e->td;
    memset(&td, 0, sizeof(td));
    td.T = T; td.A = A; td.n_sims = n; td.players = c.num_players;
    td.slots_per_tree = 1 + A + n * A;
    const size_t S = (size_t)T * td.slots_per_tree;
    MZB_CUDA(e->meta.alloc(S * sizeof(int4)));
    MZB_CUDA(e->prior.alloc(S * sizeof(double)));
    MZB_CUDA(e->value_sum.alloc(S * sizeof(double)));
    MZB_CUDA(e->reward.alloc(S * sizeof(double)));
    MZB_CUDA(e->mm_min.alloc(T * sizeof(double)));
    MZB_CUDA(e->mm_max.alloc(T * sizeof(double)));
    MZB_CUDA(e->root_to_play.alloc(T * sizeof(int)));
    MZB_CUDA(e->n_legal.alloc(T * sizeof(int)));
    MZB_CUDA(e->n_slots.alloc(T * sizeof(int)));
    MZB_CUDA(e->max_depth.alloc(T * sizeof(int)));
    MZB_CUDA(e->mt.alloc((size_t)T * 624 * sizeof(uint32_t)));
    MZB_CUDA(e->mt_idx.alloc(T * sizeof(int)));
    MZB_CUDA(e->path_slot.alloc((size_t)T * (n + 2) * sizeof(int)));
    MZB_CUDA(e->path_len.alloc(T * sizeof(int)));
    MZB_CUDA(e->in_slot.alloc(T * sizeof(int)));
    MZB_CUDA(e->out_slot.alloc(T * sizeof(int)));
    MZB_CUDA(e->action.alloc(T * sizeof(int)));
    MZB_CUDA(e->value.alloc(T * sizeof(double)));
    MZB_CUDA(e->reward_out.alloc(T * sizeof(double)));
    MZB_CUDA(e->policy.alloc((size_t)T * A * sizeof(double)));
    MZB_CUDA(e->pbc_table.alloc(2 * (n + 2) * sizeof(double)));
    td.meta = e->meta.as<int4>(); td.prior = e->prior.as<double>(); td.value_sum = e->value_sum.as<double>();
    td.reward = e->reward.as<double>(); td.mm_min = e->mm_min.as<double>(); td.mm_max = e->mm_max.as<double>();
    td.root_to_play = e->root_to_play.as<int>(); td.n_legal = e->n_legal.as<int>(); td.n_slots = e->n_slots.as<int>();
    td.max_depth = e->max_depth.as<int>(); td.mt = e->mt.as<uint32_t>(); td.mt_idx = e->mt_idx.as<int>();
    td.path_slot = e->path_slot.as<int>(); td.path_len = e->path_len.as<int>();
    td.in_slot = e->in_slot.as<int>(); td.out_slot = e->out_slot.as<int>(); td.action = e->action.as<int>();
    td.value = e->value.as<double>(); td.reward_out = e->reward_out.as<double>(); td.policy = e->policy.as<double>();
    td.discount = c.discount;
    td.init_min = c.value_lowerbound;   // MinMaxStats(minimum = value_lowerbound, maximum = value_upperbound), mcts.cpp:25
    td.init_max = c.value_upperbound;
    td.dirichlet_epsilon = c.dirichlet_epsilon;
    // pb_c table: std::log is glibc's on the reference side, so evaluate it on the host with the same libm
    // second half: std::sqrt(N) (mcts.cpp:130), correctly rounded on host and device alike
    std::vector<double> tab(2 * (n + 2));
    for (int N = 0; N < n + 2; ++N) tab[N] = std::log((N + c.pb_c_base + 1) / c.pb_c_base) + c.pb_c_init;  // mcts.cpp:129
    for (int N = 0; N < n + 2; ++N) tab[n + 2 + N] = std::sqrt((double)N);
    MZB_CUDA(cudaMemcpy(e->pbc_table.p, tab.data(), tab.size() * sizeof(double), cudaMemcpyHostToDevice));
    td.pbc_table = e->pbc_table.as<double>();
    // inputs / results
    MZB_CUDA(e->d_legal.alloc(T * sizeof(uint32_t)));
    MZB_CUDA(e->d_to_play.alloc(T * sizeof(int)));
    MZB_CUDA(e->d_add_noise.alloc(T));
    MZB_CUDA(e->d_noise.alloc((size_t)T * A * sizeof(double)));
    MZB_CUDA(e->d_root_slots.alloc(T * sizeof(int)));
    MZB_CUDA(e->d_iota.alloc((size_t)2 * T * sizeof(int)));
    MZB_CUDA(e->r_root_value.alloc(T * sizeof(double)));
    MZB_CUDA(e->r_depth.alloc(T * sizeof(int)));
    MZB_CUDA(e->r_nchild.alloc(T * sizeof(int)));
    MZB_CUDA(e->r_actions.alloc((size_t)T * A * sizeof(int)));
    MZB_CUDA(e->r_visits.alloc((size_t)T * A * sizeof(int)));
    MZB_CUDA(e->r_rel.alloc((size_t)T * A * sizeof(double)));
    MZB_CUDA(e->r_policy.alloc((size_t)T * A * sizeof(double)));
    std::vector<int> roots(T), iota(2 * T);
    for (int t = 0; t < T; ++t) roots[t] = t * (n + 1);
    for (int i = 0; i < 2 * T; ++i) iota[i] = i;
    MZB_CUDA(cudaMemcpy(e->d_root_slots.p, roots.data(), T * sizeof(int), cudaMemcpyHostToDevice));
    MZB_CUDA(cudaMemcpy(e->d_iota.p, iota.data(), 2 * T * sizeof(int), cudaMemcpyHostToDevice));
    // seeds: slot i <- seed + i
    std::vector<int> seeds(T);
    for (int t = 0; t < T; ++t) seeds[t] = c.seed + t;
    MZB_CUDA(cudaMemcpy(e->d_to_play.p, seeds.data(), T * sizeof(int), cudaMemcpyHostToDevice));
    launch_mt_seed(td.mt, td.mt_idx, e->d_to_play.as<int>(), 0, T, e->stream);
    MZB_CUDA(cudaStreamSynchronize(e->stream));
    return MZB_OK;
}

int set_recording_buffers(mzb_engine* e, bool on) {
    TreeDev& td = e->td;
    if (on) {
        const size_t T = td.T, n = td.n_sims, A = td.A;
        if (!e->trace_parent.p) {
            MZB_CUDA(e->trace_parent.alloc(T * n * sizeof(int)));
            MZB_CUDA(e->trace_action.alloc(T * n * sizeof(int)));
            MZB_CUDA(e->rec_value.alloc(T * (n + 1) * sizeof(double)));
            MZB_CUDA(e->rec_reward.alloc(T * (n + 1) * sizeof(double)));
            MZB_CUDA(e->rec_policy.alloc(T * (n + 1) * A * sizeof(double)));
        }
        td.trace_parent = e->trace_parent.as<int>(); td.trace_action = e->trace_action.as<int>();
        td.rec_value = e->rec_value.as<double>(); td.rec_reward = e->rec_reward.as<double>();
        td.rec_policy = e->rec_policy.as<double>();
    } else {
        td.trace_parent = td.trace_action = nullptr;
        td.rec_value = td.rec_reward = td.rec_policy = nullptr;
    }
    e->recording = on;
    return MZB_OK;
}

}  // namespace

extern "C" void mzb_nccl_release(mzb_engine* e);   // engine_nccl.cu

extern "C" {

const char* mzb_last_error(void) { return mzb::get_error(); }
const char* mzb_version(void) { return "mzb200 0.1 (sm_100a)"; }

int mzb_engine_create(const mzb_config* config, int device, mzb_engine** out) {
    MZB_REQUIRE(config && out, "null argument");
    MZB_REQUIRE(config->max_trees >= 1, "max_trees must be >= 1");
    MZB_REQUIRE(config->num_simulations >= 0, "num_simulations must be >= 0");
    // every early return below hands the half-built engine to mzb_engine_destroy (streams, events, buffers, communicator)
    struct Destroy { void operator()(mzb_engine* p) const { mzb_engine_destroy(p); } };
    std::unique_ptr<mzb_engine, Destroy> e(new mzb_engine());
    e->cfg = *config;
    e->device = device;
    memset(&e->timing, 0, sizeof(e->timing));
    memset(&e->se, 0, sizeof(e->se));
    int rc = derive_dims(e->cfg, &e->dims, &e->venc, &e->renc);
    if (rc) return rc;
    if (config->action_planes) {
        size_t n = (size_t)e->dims.A * e->dims.ach * e->dims.hw();
        e->action_planes.assign(config->action_planes, config->action_planes + n);
        e->cfg.action_planes = e->action_planes.data();
    }
    int ndev = 0;
    MZB_CUDA(cudaGetDeviceCount(&ndev));
    MZB_REQUIRE(device >= 0 && device < ndev, "CUDA device %d not present (%d devices)", device, ndev);
    MZB_CUDA(cudaSetDevice(device));
    {
        // tree kernels alone (SYNTHETIC): the search stream outranks the state mover's, so the latency-bound tree step of simulation
        // i + 1 gets its few blocks onto the SMs between the mover's thousands of copy CTAs instead of queueing behind them
        int lowest = 0, greatest = 0;
        MZB_CUDA(cudaDeviceGetStreamPriorityRange(&lowest, &greatest));
        MZB_CUDA(cudaStreamCreateWithPriority(&e->stream, cudaStreamNonBlocking, is_synthetic(e->cfg.evaluator) ? greatest : lowest));
    }
    MZB_CUDA(cudaStreamCreateWithFlags(&e->aux, cudaStreamNonBlocking));
    for (int i = 0; i < mzb_engine::MAX_PARTS - 1; ++i) MZB_CUDA(cudaStreamCreateWithFlags(&e->pipe[i], cudaStreamNonBlocking));
    MZB_CUDA(cudaEventCreate(&e->ev_start));
    MZB_CUDA(cudaEventCreate(&e->ev_root));
    MZB_CUDA(cudaEventCreate(&e->ev_end));
    MZB_CUDA(cudaEventCreateWithFlags(&e->ev_aux, cudaEventDisableTiming));
    MZB_CUDA(cudaEventCreateWithFlags(&e->ev_fork, cudaEventDisableTiming));
    for (int i = 0; i < mzb_engine::MAX_PARTS - 1; ++i) MZB_CUDA(cudaEventCreateWithFlags(&e->ev_join[i], cudaEventDisableTiming));
    for (int i = 0; i < mzb_engine::MAX_PARTS; ++i) MZB_CUDA(cudaEventCreateWithFlags(&e->ev_tower[i], cudaEventDisableTiming));
    rc = alloc_trees(e.get());
    if (rc) return rc;
    const int T = e->cfg.max_trees, n = e->cfg.num_simulations;
    if (is_synthetic(e->cfg.evaluator)) {
        e->se.slot_floats = e->dims.hidden_size();
        e->se.vlo = e->cfg.se_value_lo; e->se.vhi = e->cfg.se_value_hi;
        e->se.rlo = e->cfg.se_reward_lo; e->se.rhi = e->cfg.se_reward_hi;
        MZB_CUDA(e->se_cid.alloc((size_t)T * (n + 1) * sizeof(uint32_t)));
        MZB_CUDA(e->se_pool.alloc((size_t)T * (n + 1) * e->se.slot_floats * sizeof(float)));
        e->se.node_cid = e->se_cid.as<uint32_t>();
        e->se.pool = e->se_pool.as<float>();
        e->se.bulk = se_move_available(e->se.slot_floats) ? 1 : 0;
        MZB_CUDA(e->se_mv.alloc((size_t)4 * T * sizeof(int)));
        for (int k = 0; k < 2; ++k) {
            MZB_CUDA(cudaEventCreateWithFlags(&e->ev_se_out[k], cudaEventDisableTiming));
            MZB_CUDA(cudaEventCreateWithFlags(&e->ev_se_mv[k], cudaEventDisableTiming));
        }
        MZB_CUDA(e->d_obs.alloc((size_t)T * (e->cfg.evaluator == MZB_EVAL_SYNTHETIC_OBS ? e->dims.initial_flat() : 1) * sizeof(float)));
    } else if (e->cfg.evaluator == MZB_EVAL_HOST) {
        MZB_CUDA(e->d_obs.alloc((size_t)T * e->dims.initial_flat() * sizeof(float)));
    } else {
        MZB_REQUIRE(e->cfg.evaluator == MZB_EVAL_NETWORK, "unknown evaluator kind %d", e->cfg.evaluator);
        e->net.reset(e->cfg.precision == MZB_BF16 ? make_bf16_backend() : make_fp32_backend());
        MZB_REQUIRE(e->net != nullptr, "precision mode %d is not available in this build", e->cfg.precision);
        MZB_CUDA(e->d_obs.alloc((size_t)T * e->dims.initial_flat() * sizeof(float)));
    }
    *out = e.release();
    return MZB_OK;
}

void mzb_engine_destroy(mzb_engine* e) {
    if (!e) return;
    cudaSetDevice(e->device);
    cudaDeviceSynchronize();
    if (e->h_pin) cudaFreeHost(e->h_pin);
    if (e->ev_start) cudaEventDestroy(e->ev_start);
    if (e->ev_root) cudaEventDestroy(e->ev_root);
    if (e->ev_end) cudaEventDestroy(e->ev_end);
    if (e->ev_aux) cudaEventDestroy(e->ev_aux);
    if (e->ev_fork) cudaEventDestroy(e->ev_fork);
    for (int i = 0; i < mzb_engine::MAX_PARTS - 1; ++i) if (e->ev_join[i]) cudaEventDestroy(e->ev_join[i]);
    for (int i = 0; i < mzb_engine::MAX_PARTS; ++i) if (e->ev_tower[i]) cudaEventDestroy(e->ev_tower[i]);
    if (e->sim_graph) cudaGraphExecDestroy(e->sim_graph);
    for (int i = 0; i < mzb_engine::MAX_PARTS; ++i) { if (e->hi[i]) cudaStreamDestroy(e->hi[i]); if (e->ev_tree[i]) cudaEventDestroy(e->ev_tree[i]); }
    for (int k = 0; k < 2; ++k) { if (e->ev_se_out[k]) cudaEventDestroy(e->ev_se_out[k]); if (e->ev_se_mv[k]) cudaEventDestroy(e->ev_se_mv[k]); }
    if (e->learn_step && e->learn_step_free) e->learn_step_free(e->learn_step);
    mzb_nccl_release(e);
    if (e->sync_stream) cudaStreamDestroy(e->sync_stream);
    if (e->ev_sync) cudaEventDestroy(e->ev_sync);
    e->net.reset();
    if (e->stream) cudaStreamDestroy(e->stream);
    if (e->aux) cudaStreamDestroy(e->aux);
    for (int i = 0; i < mzb_engine::MAX_PARTS - 1; ++i) if (e->pipe[i]) cudaStreamDestroy(e->pipe[i]);
    delete e;
}

int mzb_get_shapes(mzb_engine* e, mzb_shapes* o) {
    MZB_REQUIRE(e && o, "null argument");
    memset(o, 0, sizeof(*o));
    o->initial_channels = e->dims.cin; o->initial_flat_size = e->dims.initial_flat();
    o->enc_c = e->dims.C; o->enc_h = e->dims.h; o->enc_w = e->dims.w; o->recurrent_flat_size = e->dims.hidden_size();
    o->value_support_size = e->dims.Sv; o->reward_support_size = e->dims.Sr;
    o->value_step = e->venc.step; o->value_min = e->venc.min_value;
    o->reward_step = e->renc.step; o->reward_min = e->renc.min_value;
    o->weight_blob_bytes = e->net ? (int64_t)e->net->blob_bytes() : 0;
    o->node_pool_bytes = e->net ? (int64_t)e->net->pool_bytes() : (int64_t)e->se_pool.bytes;
    return MZB_OK;
}

int mzb_load_weights(mzb_engine* e, const mzb_tensor_desc* tensors, int n_tensors) {
    MZB_REQUIRE(e && tensors && n_tensors > 0, "null argument");
    MZB_REQUIRE(e->net != nullptr, "engine was created with the synthetic or a host evaluator; it has no network");
    MZB_CUDA(cudaSetDevice(e->device));
    TensorMap tm;
    for (int i = 0; i < n_tensors; ++i) {
        MZB_REQUIRE(tensors[i].name && tensors[i].data && tensors[i].ndim >= 0 && tensors[i].ndim <= 4, "bad tensor descriptor %d", i);
        std::vector<int64_t> dims(tensors[i].dims, tensors[i].dims + tensors[i].ndim);
        tm[tensors[i].name] = std::make_pair(tensors[i].data, dims);
    }
    int rc = build_plan(e->cfg, tm, &e->plan);
    if (rc) return rc;
    MZB_CUDA(cudaStreamSynchronize(e->stream));
    rc = e->net->upload(e->plan);
    if (rc) return rc;
    if (!e->weights_loaded) {
        rc = e->net->alloc_pool((size_t)e->cfg.max_trees * (e->cfg.num_simulations + 1) + 2 * (size_t)e->cfg.max_trees, e->cfg.max_trees);
        if (rc) return rc;
    }
    e->weights_loaded = true;
    e->raw_weights.clear();
    e->raw_longs.clear();
    ++e->weights_serial;
    for (int i = 0; i < n_tensors; ++i) {
        mzb_engine::RawTensor r;
        r.name = tensors[i].name;
        r.dims.assign(tensors[i].dims, tensors[i].dims + tensors[i].ndim);
        size_t numel = 1;
        for (int64_t d : r.dims) numel *= (size_t)d;
        r.data.assign(tensors[i].data, tensors[i].data + numel);
        e->raw_weights.push_back(std::move(r));
    }
    return MZB_OK;
}

}  // extern "C" (checkpoint.cpp's view of the engine)
namespace mzb {
void engine_set_long_scalars(mzb_engine* e, const std::vector<std::pair<std::string, int64_t>>& longs) { e->raw_longs = longs; }
int engine_named_tensors(mzb_engine* e, std::vector<mzb_tensor_desc>* tensors, std::vector<const char*>* long_names, std::vector<int64_t>* long_values) {
    MZB_REQUIRE(e->weights_loaded && !e->raw_weights.empty(), "no weights loaded (mzb_load_weights / mzb_load_checkpoint): nothing to save");
    for (const mzb_engine::RawTensor& r : e->raw_weights) {
        mzb_tensor_desc d;
        memset(&d, 0, sizeof(d));
        d.name = r.name.c_str(); d.data = r.data.data(); d.ndim = (int32_t)r.dims.size();
        for (size_t k = 0; k < r.dims.size(); ++k) d.dims[k] = r.dims[k];
        tensors->push_back(d);
    }
    for (const auto& kv : e->raw_longs) { long_names->push_back(kv.first.c_str()); long_values->push_back(kv.second); }
    return MZB_OK;
}
}  // namespace mzb
extern "C" {

int mzb_weights_checksum(mzb_engine* e, uint64_t* out) {
    MZB_REQUIRE(e && out && e->net && e->weights_loaded, "no weights loaded");
    MZB_CUDA(cudaSetDevice(e->device));
    std::vector<uint8_t> h(e->net->blob_bytes());
    MZB_CUDA(cudaStreamSynchronize(e->stream));   // a commit or a load queued on the search stream lands first
    MZB_CUDA(cudaMemcpy(h.data(), e->net->blob_ptr(), h.size(), cudaMemcpyDeviceToHost));
    uint64_t x = 1469598103934665603ULL;
    for (uint8_t b : h) { x ^= b; x *= 1099511628211ULL; }
    *out = x;
    return MZB_OK;
}

int mzb_search_seed(mzb_engine* e, int first_slot, int n, const int32_t* seeds) {
    MZB_REQUIRE(e && seeds && first_slot >= 0 && n >= 0 && first_slot + n <= e->cfg.max_trees, "bad slot range");
    MZB_CUDA(cudaSetDevice(e->device));
    MZB_CUDA(cudaMemcpyAsync(e->d_to_play.p, seeds, n * sizeof(int), cudaMemcpyHostToDevice, e->stream));
    launch_mt_seed(e->td.mt, e->td.mt_idx, e->d_to_play.as<int>(), first_slot, n, e->stream);
    MZB_CUDA(cudaGetLastError());
    MZB_CUDA(cudaStreamSynchronize(e->stream));
    return MZB_OK;
}

int mzb_set_host_evaluator(mzb_engine* e, const mzb_host_evaluator* evaluator) {
    MZB_REQUIRE(e && evaluator && evaluator->initial && evaluator->recurrent, "null argument");
    MZB_REQUIRE(e->cfg.evaluator == MZB_EVAL_HOST, "engine was not created with MZB_EVAL_HOST");
    e->host_eval = *evaluator;
    return MZB_OK;
}

// ---- unit-level parity hooks (tests/mcts_test.cpp's hand-built trees on the device step)
static int debug_slot_check(mzb_engine* e, int slot) {
    MZB_REQUIRE(e, "null argument");
    MZB_REQUIRE(slot >= 0 && slot < e->cfg.max_trees, "slot %d outside [0, %d)", slot, e->cfg.max_trees);
    MZB_CUDA(cudaSetDevice(e->device));
    return MZB_OK;
}

int mzb_debug_tree_write(mzb_engine* e, int slot, const mzb_debug_tree* t) {
    int rc = debug_slot_check(e, slot);
    if (rc) return rc;
    MZB_REQUIRE(t && t->meta && t->prior && t->value_sum && t->reward, "null argument");
    const TreeDev& td = e->td;
    MZB_REQUIRE(t->n_nodes >= 1 && t->n_nodes + td.A <= td.slots_per_tree, "%d nodes do not fit a tree slot (%d node slots)", t->n_nodes, td.slots_per_tree);
    MZB_REQUIRE(t->n_root_children >= 0 && t->n_root_children <= td.A, "bad root child count");
    MZB_REQUIRE(t->root_to_play >= 0 && t->root_to_play < td.players, "root_to_play out of range");
    for (int i = 0; i < t->n_nodes; ++i) {
        const int cb = t->meta[4 * i + 2];
        MZB_REQUIRE(cb < 0 || (cb > i && cb + (i == 0 ? t->n_root_children : td.A) <= t->n_nodes), "node %d: child_base %d outside the tree", i, cb);
    }
    const size_t sb = (size_t)slot * td.slots_per_tree, n = t->n_nodes;
    MZB_CUDA(cudaMemcpyAsync(td.meta + sb, t->meta, n * sizeof(int4), cudaMemcpyHostToDevice, e->stream));
    MZB_CUDA(cudaMemcpyAsync(td.prior + sb, t->prior, n * sizeof(double), cudaMemcpyHostToDevice, e->stream));
    MZB_CUDA(cudaMemcpyAsync(td.value_sum + sb, t->value_sum, n * sizeof(double), cudaMemcpyHostToDevice, e->stream));
    MZB_CUDA(cudaMemcpyAsync(td.reward + sb, t->reward, n * sizeof(double), cudaMemcpyHostToDevice, e->stream));
    MZB_CUDA(cudaMemcpyAsync(td.mm_min + slot, &t->minmax_minimum, sizeof(double), cudaMemcpyHostToDevice, e->stream));
    MZB_CUDA(cudaMemcpyAsync(td.mm_max + slot, &t->minmax_maximum, sizeof(double), cudaMemcpyHostToDevice, e->stream));
    const int zero = 0;
    MZB_CUDA(cudaMemcpyAsync(td.root_to_play + slot, &t->root_to_play, sizeof(int), cudaMemcpyHostToDevice, e->stream));
    MZB_CUDA(cudaMemcpyAsync(td.n_legal + slot, &t->n_root_children, sizeof(int), cudaMemcpyHostToDevice, e->stream));
    MZB_CUDA(cudaMemcpyAsync(td.n_slots + slot, &t->n_nodes, sizeof(int), cudaMemcpyHostToDevice, e->stream));
    MZB_CUDA(cudaMemcpyAsync(td.max_depth + slot, &zero, sizeof(int), cudaMemcpyHostToDevice, e->stream));
    MZB_CUDA(cudaStreamSynchronize(e->stream));   // the uploads ride the engine's stream: ordered before the step that follows
    return MZB_OK;
}

int mzb_debug_tree_read(mzb_engine* e, int slot, int n_nodes, int32_t* meta, double* prior, double* value_sum, double* reward, double* minmax) {
    int rc = debug_slot_check(e, slot);
    if (rc) return rc;
    const TreeDev& td = e->td;
    MZB_REQUIRE(n_nodes >= 1 && n_nodes <= td.slots_per_tree, "bad node count");
    const size_t sb = (size_t)slot * td.slots_per_tree, n = n_nodes;
    if (meta) MZB_CUDA(cudaMemcpy(meta, td.meta + sb, n * sizeof(int4), cudaMemcpyDeviceToHost));
    if (prior) MZB_CUDA(cudaMemcpy(prior, td.prior + sb, n * sizeof(double), cudaMemcpyDeviceToHost));
    if (value_sum) MZB_CUDA(cudaMemcpy(value_sum, td.value_sum + sb, n * sizeof(double), cudaMemcpyDeviceToHost));
    if (reward) MZB_CUDA(cudaMemcpy(reward, td.reward + sb, n * sizeof(double), cudaMemcpyDeviceToHost));
    if (minmax) {
        MZB_CUDA(cudaMemcpy(minmax, td.mm_min + slot, sizeof(double), cudaMemcpyDeviceToHost));
        MZB_CUDA(cudaMemcpy(minmax + 1, td.mm_max + slot, sizeof(double), cudaMemcpyDeviceToHost));
    }
    return MZB_OK;
}

int mzb_debug_select_child(mzb_engine* e, int slot, int restore_rng, int32_t* chosen_node, double* scores) {
    int rc = debug_slot_check(e, slot);
    if (rc) return rc;
    MZB_REQUIRE(chosen_node, "null argument");
    MZB_REQUIRE(e->cfg.num_simulations >= 1, "needs num_simulations >= 1");
    if (!e->dbg_score.p) MZB_CUDA(e->dbg_score.alloc(32 * sizeof(double)));
    int4 root;
    MZB_CUDA(cudaMemcpy(&root, e->td.meta + (size_t)slot * e->td.slots_per_tree, sizeof(int4), cudaMemcpyDeviceToHost));
    MZB_REQUIRE(root.z >= 0, "the root is not expanded");
    MZB_REQUIRE(root.x >= 0 && root.x <= e->cfg.num_simulations + 1, "root visit count %d outside the pb_c table", root.x);
    TreeDev td = e->td;
    td.dbg_score = e->dbg_score.as<double>();
    td.trace_parent = nullptr; td.trace_action = nullptr; td.rec_value = nullptr; td.rec_reward = nullptr; td.rec_policy = nullptr;
    std::vector<uint32_t> saved_mt(restore_rng ? 624 : 0);
    int saved_idx = 0;
    if (restore_rng) {
        MZB_CUDA(cudaMemcpy(saved_mt.data(), e->td.mt + (size_t)slot * 624, 624 * sizeof(uint32_t), cudaMemcpyDeviceToHost));
        MZB_CUDA(cudaMemcpy(&saved_idx, e->td.mt_idx + slot, sizeof(int), cudaMemcpyDeviceToHost));
    }
    // step 0 = selection only; stops below the root when the chosen child is not expanded
    launch_tree_step_range(td, slot, slot + 1, 0, e->stream);
    MZB_CUDA(cudaGetLastError());
    MZB_CUDA(cudaStreamSynchronize(e->stream));
    if (restore_rng) {
        MZB_CUDA(cudaMemcpyAsync(e->td.mt + (size_t)slot * 624, saved_mt.data(), 624 * sizeof(uint32_t), cudaMemcpyHostToDevice, e->stream));
        MZB_CUDA(cudaMemcpyAsync(e->td.mt_idx + slot, &saved_idx, sizeof(int), cudaMemcpyHostToDevice, e->stream));
        MZB_CUDA(cudaStreamSynchronize(e->stream));
    }
    std::vector<int> path(e->cfg.num_simulations + 2);
    MZB_CUDA(cudaMemcpy(path.data(), e->td.path_slot + (size_t)slot * (e->cfg.num_simulations + 2), 2 * sizeof(int), cudaMemcpyDeviceToHost));
    *chosen_node = path[1];
    if (scores) MZB_CUDA(cudaMemcpy(scores, e->dbg_score.p, e->cfg.num_actions * sizeof(double), cudaMemcpyDeviceToHost));
    return MZB_OK;
}

int mzb_debug_backpropagate(mzb_engine* e, int slot, const int32_t* path, int len, double value, double leaf_reward, const double* policy) {
    int rc = debug_slot_check(e, slot);
    if (rc) return rc;
    const TreeDev& td0 = e->td;
    const int n = e->cfg.num_simulations, A = e->cfg.num_actions;
    MZB_REQUIRE(path && policy, "null argument");
    MZB_REQUIRE(n >= 1 && len >= 1 && len <= n + 2, "path length %d outside [1, num_simulations + 2]", len);
    int n_slots = 0;
    MZB_CUDA(cudaMemcpy(&n_slots, td0.n_slots + slot, sizeof(int), cudaMemcpyDeviceToHost));
    MZB_REQUIRE(n_slots + A <= td0.slots_per_tree, "no room to expand the leaf");
    for (int i = 0; i < len; ++i) MZB_REQUIRE(path[i] >= 0 && path[i] < n_slots, "path[%d] = %d is not a node", i, path[i]);
    MZB_CUDA(cudaMemcpyAsync(td0.path_slot + (size_t)slot * (n + 2), path, len * sizeof(int), cudaMemcpyHostToDevice, e->stream));
    MZB_CUDA(cudaMemcpyAsync(td0.path_len + slot, &len, sizeof(int), cudaMemcpyHostToDevice, e->stream));
    MZB_CUDA(cudaMemcpyAsync(td0.value + slot, &value, sizeof(double), cudaMemcpyHostToDevice, e->stream));
    MZB_CUDA(cudaMemcpyAsync(td0.reward_out + slot, &leaf_reward, sizeof(double), cudaMemcpyHostToDevice, e->stream));
    MZB_CUDA(cudaMemcpyAsync(td0.policy + (size_t)slot * A, policy, A * sizeof(double), cudaMemcpyHostToDevice, e->stream));
    TreeDev td = e->td;
    td.trace_parent = nullptr; td.trace_action = nullptr; td.rec_value = nullptr; td.rec_reward = nullptr; td.rec_policy = nullptr;
    // step n = expansion + backup only (no further selection)
    launch_tree_step_range(td, slot, slot + 1, n, e->stream);
    MZB_CUDA(cudaGetLastError());
    MZB_CUDA(cudaStreamSynchronize(e->stream));
    return MZB_OK;
}

int mzb_debug_se_state(mzb_engine* e, int tree, int node, float* out) {
    int rc = debug_slot_check(e, tree);
    if (rc) return rc;
    MZB_REQUIRE(is_synthetic(e->cfg.evaluator) && e->se.pool, "engine was not created with the synthetic evaluator");
    MZB_REQUIRE(out && node >= 0 && node <= e->cfg.num_simulations, "node %d outside [0, num_simulations]", node);
    MZB_CUDA(cudaStreamSynchronize(e->stream));
    const size_t slot = (size_t)tree * (e->cfg.num_simulations + 1) + node;
    MZB_CUDA(cudaMemcpy(out, e->se.pool + slot * e->se.slot_floats, (size_t)e->se.slot_floats * sizeof(float), cudaMemcpyDeviceToHost));
    return MZB_OK;
}

int mzb_search_set_recording(mzb_engine* e, int enabled) {
    MZB_REQUIRE(e, "null argument");
    MZB_CUDA(cudaSetDevice(e->device));
    return set_recording_buffers(e, enabled != 0);
}

}  // extern "C"
