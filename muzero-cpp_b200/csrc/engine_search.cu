// engine_search.cu — the lock-step batched search (T x MCTS::run, mcts.cpp:34-87) and the batch inference entry
// points of the C ABI.  One host thread enqueues everything on one stream; the only host work inside a search is
// the Dirichlet noise (std::gamma_distribution on the slots' own mt19937 streams, util.cpp:28-43), overlapped with
// the initial inference.
#include <cstring>
#include <numeric>
#include <random>
#include <thread>

#include "engine.h"

using namespace mzb;

namespace {

// util::sample_dirichlet (util.cpp:28-43) on a tree slot's RNG stream
void sample_dirichlet(HostMt& rng, double alpha, int n, double* out) {
    std::gamma_distribution<double> gamma_dist(alpha, 1.0);
    for (int i = 0; i < n; ++i) out[i] = gamma_dist(rng);
    double sum = std::reduce(out, out + n);
    for (int i = 0; i < n; ++i) out[i] /= sum;
}

int popcount32(uint32_t x) { return __builtin_popcount(x); }

NetOutputs tree_outputs(mzb_engine* e) {
    NetOutputs o;
    o.value = e->td.value; o.reward = e->td.reward_out; o.policy = e->td.policy;
    return o;
}

// d_obs: device rows.  Everything else host.
int search_impl(mzb_engine* e, const float* d_obs, int obs_stride, const uint32_t* legal_mask, const int32_t* to_play,
                const uint8_t* add_noise, int T, mzb_search_result* out) {
    const mzb_config& c = e->cfg;
    const int A = c.num_actions, n = c.num_simulations;
    TreeDev& td = e->td;
    const bool synthetic = is_synthetic(c.evaluator);
    const bool host = c.evaluator == MZB_EVAL_HOST;
    const uint32_t full = (A >= 32) ? 0xffffffffu : ((1u << A) - 1u);
    bool any_noise = false;
    for (int t = 0; t < T; ++t) {
        MZB_REQUIRE((legal_mask[t] & full) != 0 && (legal_mask[t] & ~full) == 0, "tree %d: legal_mask must select 1..A actions (mcts.cpp:37)", t);
        MZB_REQUIRE(to_play[t] >= 0 && to_play[t] < c.num_players, "tree %d: to_play out of range", t);
        any_noise = any_noise || add_noise[t];
    }
    cudaStream_t s = e->stream;
    int64_t launches = 0;
    MZB_CUDA(cudaEventRecord(e->ev_start, s));
    // ---- root: initial inference (async) ----
    std::vector<int32_t> hv_slot, hv_parent, hv_action;
    std::vector<double> hv_out;
    // MZB_EVAL_HOST: outputs of one Evaluator call for all rows -> the arrays the tree kernels read
    auto host_outputs_to_device = [&]() -> int {
        MZB_CUDA(cudaMemcpyAsync(td.value, hv_out.data(), T * sizeof(double), cudaMemcpyHostToDevice, s));
        MZB_CUDA(cudaMemcpyAsync(td.reward_out, hv_out.data() + T, T * sizeof(double), cudaMemcpyHostToDevice, s));
        MZB_CUDA(cudaMemcpyAsync(td.policy, hv_out.data() + 2 * (size_t)T, (size_t)T * A * sizeof(double), cudaMemcpyHostToDevice, s));
        MZB_CUDA(cudaStreamSynchronize(s));
        return MZB_OK;
    };
    if (host) {
        MZB_REQUIRE(e->host_eval.initial && e->host_eval.recurrent, "no host evaluator set (mzb_set_host_evaluator)");
        MZB_REQUIRE(e->h_obs_rows, "a host evaluator needs host observations (mzb_search_batch / mzb_search_batch_slots)");
        hv_slot.resize(T); hv_parent.resize(T); hv_action.resize(T);
        hv_out.assign((size_t)T * (2 + A), 0.0);
        for (int t = 0; t < T; ++t) hv_slot[t] = e->h_slot_rows ? e->h_slot_rows[t] : t;
        int hrc = e->host_eval.initial(e->host_eval.user, T, hv_slot.data(), e->h_obs_rows, obs_stride, hv_out.data(),
                                       hv_out.data() + T, hv_out.data() + 2 * (size_t)T);
        MZB_REQUIRE(hrc == 0, "host evaluator: InitialInference failed (%d)", hrc);
        hrc = host_outputs_to_device();
        if (hrc) return hrc;
    } else if (synthetic) {
        launch_se_initial(td, e->se, T, d_obs, obs_stride, s);
        ++launches;
    } else {
        int rc = e->net->initial(d_obs, T, e->d_root_slots.as<int>(), tree_outputs(e), s);
        if (rc) return rc;
        launches += e->net->launches_initial();
    }
    // ---- host side, overlapped: per-tree inputs and Dirichlet noise ----
    size_t pin_need = (size_t)T * (sizeof(uint32_t) + sizeof(int) + 1 + (size_t)A * sizeof(double) + 624 * sizeof(uint32_t) + sizeof(int)) + 256;
    int rc = ensure_pinned(e, pin_need);
    if (rc) return rc;
    uint8_t* pin = (uint8_t*)e->h_pin;
    double* h_noise = (double*)pin; pin += (size_t)T * A * sizeof(double);
    uint32_t* h_mt = (uint32_t*)pin; pin += (size_t)T * 624 * sizeof(uint32_t);
    uint32_t* h_legal = (uint32_t*)pin; pin += (size_t)T * sizeof(uint32_t);
    int* h_to_play = (int*)pin; pin += (size_t)T * sizeof(int);
    int* h_idx = (int*)pin; pin += (size_t)T * sizeof(int);
    uint8_t* h_add = pin;
    memcpy(h_legal, legal_mask, T * sizeof(uint32_t));
    memcpy(h_to_play, to_play, T * sizeof(int));
    memcpy(h_add, add_noise, T);
    cudaStream_t ax = e->aux;
    MZB_CUDA(cudaMemcpyAsync(e->d_legal.p, h_legal, T * sizeof(uint32_t), cudaMemcpyHostToDevice, ax));
    MZB_CUDA(cudaMemcpyAsync(e->d_to_play.p, h_to_play, T * sizeof(int), cudaMemcpyHostToDevice, ax));
    MZB_CUDA(cudaMemcpyAsync(e->d_add_noise.p, h_add, T, cudaMemcpyHostToDevice, ax));
    if (any_noise) {
        MZB_CUDA(cudaMemcpyAsync(h_mt, td.mt, (size_t)T * 624 * sizeof(uint32_t), cudaMemcpyDeviceToHost, ax));
        MZB_CUDA(cudaMemcpyAsync(h_idx, td.mt_idx, T * sizeof(int), cudaMemcpyDeviceToHost, ax));
        MZB_CUDA(cudaStreamSynchronize(ax));
        // every tree draws from its own generator, so the trees split over a few host threads (1024 C4 trees: ~0.5 ms of
        // gamma draws on one core, longer than the initial inference they are meant to hide behind)
        auto draw = [&](int t_lo, int t_hi) {
            for (int t = t_lo; t < t_hi; ++t) {
                if (!add_noise[t]) continue;
                HostMt rng{h_mt + (size_t)t * 624, h_idx + t};
                sample_dirichlet(rng, c.dirichlet_alpha, popcount32(legal_mask[t]), h_noise + (size_t)t * A);
            }
        };
        static const int noise_threads = getenv("MZB_NOISE_THREADS") ? std::max(1, atoi(getenv("MZB_NOISE_THREADS"))) : 4;
        const int nthr = std::min(noise_threads, std::max(1, T / 128));
        if (nthr <= 1) draw(0, T);
        else {
            std::vector<std::thread> pool;
            for (int k = 1; k < nthr; ++k) pool.emplace_back(draw, (int)((long)T * k / nthr), (int)((long)T * (k + 1) / nthr));
            draw(0, T / nthr);
            for (std::thread& th : pool) th.join();
        }
        MZB_CUDA(cudaMemcpyAsync(td.mt, h_mt, (size_t)T * 624 * sizeof(uint32_t), cudaMemcpyHostToDevice, ax));
        MZB_CUDA(cudaMemcpyAsync(td.mt_idx, h_idx, T * sizeof(int), cudaMemcpyHostToDevice, ax));
        MZB_CUDA(cudaMemcpyAsync(e->d_noise.p, h_noise, (size_t)T * A * sizeof(double), cudaMemcpyHostToDevice, ax));
    }
    MZB_CUDA(cudaEventRecord(e->ev_aux, ax));
    MZB_CUDA(cudaStreamWaitEvent(s, e->ev_aux, 0));
    launch_tree_root(td, T, e->d_legal.as<uint32_t>(), e->d_to_play.as<int>(), e->d_add_noise.as<uint8_t>(), e->d_noise.as<double>(), s);
    ++launches;
    MZB_CUDA(cudaEventRecord(e->ev_root, s));
    // ---- simulation loop: tree step <-> recurrent inference, lock-step over all trees (mcts.cpp:52-79) ----
    // Pipelined loop: the cohort is cut into K parts (default 2), each on its own stream, when the backend allows it.  Trees
    // are independent, so while one part's towers own the SMs another part's tree step (a few blocks without shared memory)
    // runs beside them and its towers queue up behind — the tree step, the launch gaps and the tail of each tower launch
    // leave the critical path.  Results are bit-identical to the single-stream loop (a tree's arithmetic does not depend on
    // its neighbours in a tile; tests/test_gpu_pipeline.py).
    static const int pipe_min = getenv("MZB_PIPELINE_MIN_TREES") ? atoi(getenv("MZB_PIPELINE_MIN_TREES")) : 256;
    static const int pipe_parts = getenv("MZB_PIPELINE_PARTS") ? std::max(2, std::min((int)mzb_engine::MAX_PARTS, atoi(getenv("MZB_PIPELINE_PARTS")))) : 2;
    const bool pipelined = !synthetic && !host && e->net->recurrent_concurrent() && pipe_min > 0 && T >= pipe_min && T >= pipe_parts;
    if (pipelined) {
        const int K = pipe_parts;
        cudaStream_t st[mzb_engine::MAX_PARTS];
        int t0[mzb_engine::MAX_PARTS + 1];
        NetOutputs o[mzb_engine::MAX_PARTS];
        e->net->pipeline_cut(T, K, t0);
        static const int alt_env = getenv("MZB_PIPELINE_ALTERNATE") ? atoi(getenv("MZB_PIPELINE_ALTERNATE")) : -1;
        const bool alternate = alt_env > 0;
        // MZB_PIPELINE_TRACE=1: a timed event behind every launch, printed as a timeline after the search (diagnostic only)
        static const bool trace = getenv("MZB_PIPELINE_TRACE") != nullptr;
        for (int k = 0; k < K; ++k) {
            st[k] = k == 0 ? s : e->pipe[k - 1];
            o[k] = tree_outputs(e);
            o[k].value += t0[k]; o[k].reward += t0[k]; o[k].policy += (size_t)t0[k] * A;
        }
        int64_t loop_launches = 0;
        static const bool tree_prio = getenv("MZB_TREE_PRIORITY") != nullptr;
        if (tree_prio && !e->hi[0]) {
            int lowest = 0, greatest = 0;
            MZB_CUDA(cudaDeviceGetStreamPriorityRange(&lowest, &greatest));
            for (int k = 0; k < mzb_engine::MAX_PARTS; ++k) {
                MZB_CUDA(cudaStreamCreateWithPriority(&e->hi[k], cudaStreamNonBlocking, greatest));
                MZB_CUDA(cudaEventCreateWithFlags(&e->ev_tree[k], cudaEventDisableTiming));
            }
        }
        // fork -> every simulation's launches on the parts' streams -> join, as one enqueue: run eagerly, or captured into a CUDA graph
        auto enqueue = [&]() -> int {
            int lrc = MZB_OK;
            loop_launches = 0;
            MZB_CUDA(cudaEventRecord(e->ev_fork, s));
            for (int k = 1; k < K; ++k) MZB_CUDA(cudaStreamWaitEvent(st[k], e->ev_fork, 0));
            e->net->set_pipelined(true);
            std::vector<cudaEvent_t> tev;
            auto mark = [&](cudaStream_t q) { if (trace) { cudaEvent_t ev; cudaEventCreate(&ev); cudaEventRecord(ev, q); tev.push_back(ev); } };
            for (int sim = 0; sim <= n && !lrc; ++sim) {
                for (int k = 0; k < K && !lrc; ++k) {
                    if (tree_prio) {
                        // EXPERIMENT (MZB_TREE_PRIORITY=1): the part's tree step on a high-priority stream of its own, chained to the
                        // part's towers by events — its few blocks get the first SM that comes free instead of queueing behind the
                        // other parts' pending tower CTAs
                        MZB_CUDA(cudaStreamWaitEvent(e->hi[k], sim == 0 ? e->ev_fork : e->ev_tower[k], 0));
                        launch_tree_step_range(td, t0[k], t0[k + 1], sim, e->hi[k]);
                        MZB_CUDA(cudaEventRecord(e->ev_tree[k], e->hi[k]));
                        MZB_CUDA(cudaStreamWaitEvent(st[k], e->ev_tree[k], 0));
                    } else
                    launch_tree_step_range(td, t0[k], t0[k + 1], sim, st[k]);
                    ++loop_launches;
                    mark(st[k]);
                    if (sim == n) continue;
                    // EXPERIMENT (MZB_PIPELINE_ALTERNATE=1, off by default): this part's tower starts only when the previous part's tower
                    // has ended, so that the kernels behind a tower (head MLPs of a wide-head network, the next tree step) run beside the
                    // other part's tower.  Measured slower everywhere (ALE 17.4 -> 21.6 ms, c4d 11.5 -> 14.4 ms per search): serialised
                    // towers expose every launch's ramp-up and drain (~20 us), which freely overlapping launches hide.
                    if (alternate) e->net->set_tower_events((sim > 0 || k > 0) ? e->ev_tower[(k + K - 1) % K] : nullptr, e->ev_tower[k]);
                    lrc = e->net->recurrent(t0[k + 1] - t0[k], td.in_slot + t0[k], td.action + t0[k], td.out_slot + t0[k], o[k], st[k]);
                    loop_launches += e->net->launches_recurrent();
                    if (tree_prio) MZB_CUDA(cudaEventRecord(e->ev_tower[k], st[k]));
                    mark(st[k]);
                }
            }
            e->net->set_pipelined(false);
            if (trace) {
                for (int k = 0; k < K; ++k) cudaStreamSynchronize(st[k]);
                fprintf(stderr, "[pipeline] parts:");
                for (int k = 0; k < K; ++k) fprintf(stderr, " %d", t0[k + 1] - t0[k]);
                fprintf(stderr, " trees; end of (tree step | tower) per part, us since the root event\n");
                size_t i = 0;
                for (int sim = 0; sim <= n && sim < (getenv("MZB_PIPELINE_TRACE_ALL") ? n + 1 : 8); ++sim) {
                    fprintf(stderr, "[pipeline] sim %2d:", sim);
                    for (int k = 0; k < K; ++k) {
                        float a = 0.f, b = 0.f;
                        cudaEventElapsedTime(&a, e->ev_root, tev[i++]);
                        if (sim < n) cudaEventElapsedTime(&b, e->ev_root, tev[i++]);
                        fprintf(stderr, "  part %d: %8.1f | %8.1f", k, a * 1e3f, b * 1e3f);
                    }
                    fprintf(stderr, "\n");
                }
                for (cudaEvent_t ev : tev) cudaEventDestroy(ev);
            }
            // join the side streams into the search stream — also when a launch failed, so that no work is left running on them
            // behind the caller's back
            for (int k = 1; k < K; ++k) {
                MZB_CUDA(cudaEventRecord(e->ev_join[k - 1], st[k]));
                MZB_CUDA(cudaStreamWaitEvent(s, e->ev_join[k - 1], 0));
            }
            return lrc;
        };
        // CUDA graph of the loop: the launch sequence of a search depends only on the cohort size, the cut and
        // the pointers baked into the kernels' arguments.  The first search with a given key runs eagerly (it also creates the
        // backend's per-batch tables), the second is captured, later ones replay the instantiated graph: ~4 n launches + 2 n
        // event operations per search leave the host thread.  MZB_NO_GRAPH=1 keeps the eager loop.
        static const bool no_graph = getenv("MZB_NO_GRAPH") != nullptr;
        const bool graphable = !no_graph && !trace && !alternate && !getenv("MZB_TILE_TRACE") && !e->graph_failed;
        uint64_t key = 1469598103934665603ull;
        auto mix = [&](const void* p, size_t nb) { const uint8_t* b = (const uint8_t*)p; for (size_t i = 0; i < nb; ++i) { key ^= b[i]; key *= 1099511628211ull; } };
        if (graphable) {
            const void* blob = e->net->blob_ptr();
            mix(&td, sizeof(td)); mix(&T, sizeof(T)); mix(&K, sizeof(K)); mix(t0, sizeof(int) * (K + 1)); mix(&blob, sizeof(blob));
            mix(&e->weights_serial, sizeof(e->weights_serial));
        }
        if (graphable && e->sim_graph && e->sim_graph_key == key) {
            MZB_CUDA(cudaGraphLaunch(e->sim_graph, s));
            launches += e->sim_graph_launches;
        } else if (graphable && e->sim_graph_warm_key == key) {
            if (e->sim_graph) { cudaGraphExecDestroy(e->sim_graph); e->sim_graph = nullptr; }
            cudaGraph_t g = nullptr;
            cudaError_t ce = cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal);
            if (ce == cudaSuccess) {
                rc = enqueue();
                ce = cudaStreamEndCapture(s, &g);
                if (ce == cudaSuccess && !rc) ce = cudaGraphInstantiate(&e->sim_graph, g, 0);
                if (g) cudaGraphDestroy(g);
            }
            if (ce != cudaSuccess || rc || !e->sim_graph) {
                // capture is an optimisation: fall back to the eager loop for good
                cudaGetLastError();
                e->graph_failed = true; e->sim_graph = nullptr; rc = MZB_OK;
                rc = enqueue();
                launches += loop_launches;
            } else {
                e->sim_graph_key = key; e->sim_graph_launches = loop_launches;
                MZB_CUDA(cudaGraphLaunch(e->sim_graph, s));
                launches += loop_launches;
            }
        } else {
            rc = enqueue();
            launches += loop_launches;
            e->sim_graph_warm_key = key;
        }
        if (rc) { cudaStreamSynchronize(s); return rc; }
    } else if (synthetic) {
        // tree kernels alone.  The stub's hidden-state move of simulation i (se_move_kernel, HBM-bound) runs on its own stream
        // beside the tree step of simulation i + 1 (latency-bound, touches no hidden state): the outputs kernel snapshots the
        // (parent, child) slot pairs into one of two buffers, which is reused only after its mover has finished.
        cudaStream_t mv = e->pipe[0];
        const int max_t = c.max_trees;
        int used[2] = {0, 0};
        for (int sim = 0; sim <= n; ++sim) {
            launch_tree_step(td, T, sim, s);
            ++launches;
            if (sim == n) break;
            const int k = sim & 1;
            SEDev se = e->se;
            se.mv_in = e->se_mv.as<int>() + (size_t)k * 2 * max_t;
            se.mv_out = se.mv_in + max_t;
            if (se.bulk && used[k]) MZB_CUDA(cudaStreamWaitEvent(s, e->ev_se_mv[k], 0));   // the mover of simulation sim - 2 has read them
            launch_se_recurrent(td, se, T, s);
            ++launches;
            if (se.bulk) {
                MZB_CUDA(cudaEventRecord(e->ev_se_out[k], s));
                MZB_CUDA(cudaStreamWaitEvent(mv, e->ev_se_out[k], 0));
                launch_se_move(se, T, mv);
                ++launches;
                MZB_CUDA(cudaEventRecord(e->ev_se_mv[k], mv));
                used[k] = 1;
            }
        }
        for (int k = 0; k < 2; ++k)
            if (used[k]) MZB_CUDA(cudaStreamWaitEvent(s, e->ev_se_mv[k], 0));   // the search ends when the last states have landed
    } else
    for (int sim = 0; sim <= n; ++sim) {
        launch_tree_step(td, T, sim, s);
        ++launches;
        if (sim == n) break;
        if (host) {
            // Evaluator::RecurrentInference(parent.encoded_state, action) for every row (mcts.cpp:69-71)
            MZB_CUDA(cudaMemcpyAsync(hv_parent.data(), td.in_slot, T * sizeof(int), cudaMemcpyDeviceToHost, s));
            MZB_CUDA(cudaMemcpyAsync(hv_action.data(), td.action, T * sizeof(int), cudaMemcpyDeviceToHost, s));
            MZB_CUDA(cudaStreamSynchronize(s));
            for (int t = 0; t < T; ++t) hv_parent[t] -= t * (n + 1);   // pool slot -> node id (expansion order)
            int hrc = e->host_eval.recurrent(e->host_eval.user, T, hv_slot.data(), hv_parent.data(), hv_action.data(), sim + 1,
                                             hv_out.data(), hv_out.data() + T, hv_out.data() + 2 * (size_t)T);
            MZB_REQUIRE(hrc == 0, "host evaluator: RecurrentInference failed (%d)", hrc);
            hrc = host_outputs_to_device();
            if (hrc) return hrc;
        } else {
            rc = e->net->recurrent(T, td.in_slot, td.action, td.out_slot, tree_outputs(e), s);
            if (rc) return rc;
            launches += e->net->launches_recurrent();
        }
    }
    ResultDev r;
    r.root_value = e->r_root_value.as<double>(); r.max_tree_depth = e->r_depth.as<int>(); r.n_children = e->r_nchild.as<int>();
    r.child_actions = e->r_actions.as<int>(); r.child_visits = e->r_visits.as<int>();
    r.relative_visit = e->r_rel.as<double>(); r.emperical_policy = e->r_policy.as<double>();
    launch_tree_stats(td, T, r, s);
    ++launches;
    MZB_CUDA(cudaGetLastError());
    if (out) {
        if (out->root_value) MZB_CUDA(cudaMemcpyAsync(out->root_value, r.root_value, T * sizeof(double), cudaMemcpyDeviceToHost, s));
        if (out->max_tree_depth) MZB_CUDA(cudaMemcpyAsync(out->max_tree_depth, r.max_tree_depth, T * sizeof(int), cudaMemcpyDeviceToHost, s));
        if (out->n_children) MZB_CUDA(cudaMemcpyAsync(out->n_children, r.n_children, T * sizeof(int), cudaMemcpyDeviceToHost, s));
        if (out->child_actions) MZB_CUDA(cudaMemcpyAsync(out->child_actions, r.child_actions, (size_t)T * A * sizeof(int), cudaMemcpyDeviceToHost, s));
        if (out->child_visits) MZB_CUDA(cudaMemcpyAsync(out->child_visits, r.child_visits, (size_t)T * A * sizeof(int), cudaMemcpyDeviceToHost, s));
        if (out->children_relative_visit) MZB_CUDA(cudaMemcpyAsync(out->children_relative_visit, r.relative_visit, (size_t)T * A * sizeof(double), cudaMemcpyDeviceToHost, s));
        if (out->emperical_policy) MZB_CUDA(cudaMemcpyAsync(out->emperical_policy, r.emperical_policy, (size_t)T * A * sizeof(double), cudaMemcpyDeviceToHost, s));
    }
    MZB_CUDA(cudaEventRecord(e->ev_end, s));
    MZB_CUDA(cudaStreamSynchronize(s));
    MZB_CUDA(cudaGetLastError());
    float total = 0.f, root = 0.f;
    MZB_CUDA(cudaEventElapsedTime(&total, e->ev_start, e->ev_end));
    MZB_CUDA(cudaEventElapsedTime(&root, e->ev_start, e->ev_root));
    e->timing.total_ms = total; e->timing.root_ms = root; e->timing.sim_loop_ms = total - root;
    e->timing.kernel_launches = launches;
    e->timing.simulations = (int64_t)T * n;
    return MZB_OK;
}

int check_search_args(mzb_engine* e, const void* obs, const uint32_t* legal_mask, const int32_t* to_play, const uint8_t* add_noise, int T) {
    MZB_REQUIRE(e && obs && legal_mask && to_play && add_noise, "null argument");
    MZB_REQUIRE(T >= 1 && T <= e->cfg.max_trees, "n_trees %d outside [1, max_trees=%d]", T, e->cfg.max_trees);
    MZB_REQUIRE(e->cfg.evaluator != MZB_EVAL_NETWORK || e->weights_loaded, "weights not loaded (mzb_load_weights)");
    return MZB_OK;
}

}  // namespace

namespace mzb {
// A search over chosen tree slots whose observation rows already sit in e->d_obs (ordered like `slots`): gathers the slots' RNG
// rows into a compact block, searches, scatters the advanced generators back.  Shared by mzb_search_batch_slots (rows uploaded
// from the host) and mzb_search_batch_ring (rows stacked on the device from the observation ring, selfplay.cu).
int search_with_device_obs_slots(mzb_engine* e, const int32_t* slots, const uint32_t* legal_mask, const int32_t* to_play,
                                 const uint8_t* add_noise, int n_trees, mzb_search_result* out) {
    int rc = check_search_args(e, e->d_obs.p, legal_mask, to_play, add_noise, n_trees);
    if (rc) return rc;
    MZB_REQUIRE(slots, "null slot list");
    const int T = e->cfg.max_trees;
    std::vector<uint8_t> seen(T, 0);
    for (int i = 0; i < n_trees; ++i) {
        MZB_REQUIRE(slots[i] >= 0 && slots[i] < T, "row %d: slot %d outside [0, %d)", i, slots[i], T);
        MZB_REQUIRE(!seen[slots[i]], "slot %d appears twice", slots[i]);
        seen[slots[i]] = 1;
    }
    if (!e->d_slots.p) {
        MZB_CUDA(e->d_slots.alloc((size_t)T * sizeof(int)));
        MZB_CUDA(e->mt_rows.alloc((size_t)T * 624 * sizeof(uint32_t)));
        MZB_CUDA(e->mt_idx_rows.alloc((size_t)T * sizeof(int)));
    }
    cudaStream_t s = e->stream;
    MZB_CUDA(cudaMemcpyAsync(e->d_slots.p, slots, (size_t)n_trees * sizeof(int), cudaMemcpyHostToDevice, s));
    uint32_t* home = e->td.mt;
    int* home_idx = e->td.mt_idx;
    launch_mt_move(home, home_idx, e->mt_rows.as<uint32_t>(), e->mt_idx_rows.as<int>(), e->d_slots.as<int>(), n_trees, 0, s);
    MZB_CUDA(cudaStreamSynchronize(s));   // the Dirichlet draws on the aux stream read the gathered rows
    const int stride = (e->cfg.evaluator == MZB_EVAL_SYNTHETIC) ? 1 : e->dims.initial_flat();
    e->td.mt = e->mt_rows.as<uint32_t>();
    e->td.mt_idx = e->mt_idx_rows.as<int>();
    rc = search_impl(e, e->d_obs.as<float>(), stride, legal_mask, to_play, add_noise, n_trees, out);
    e->td.mt = home;
    e->td.mt_idx = home_idx;
    if (rc) return rc;
    launch_mt_move(home, home_idx, e->mt_rows.as<uint32_t>(), e->mt_idx_rows.as<int>(), e->d_slots.as<int>(), n_trees, 1, s);
    MZB_CUDA(cudaStreamSynchronize(s));
    MZB_CUDA(cudaGetLastError());
    return MZB_OK;
}
}  // namespace mzb

extern "C" {

int mzb_search_batch(mzb_engine* e, const float* obs, const uint32_t* legal_mask, const int32_t* to_play,
                     const uint8_t* add_noise, int n_trees, mzb_search_result* out) {
    int rc = check_search_args(e, obs, legal_mask, to_play, add_noise, n_trees);
    if (rc) return rc;
    MZB_CUDA(cudaSetDevice(e->device));
    const int stride = (e->cfg.evaluator == MZB_EVAL_SYNTHETIC) ? 1 : e->dims.initial_flat();
    // host -> device copy of the observations (pageable or pinned caller memory) on the search stream
    MZB_CUDA(cudaMemcpyAsync(e->d_obs.p, obs, (size_t)n_trees * stride * sizeof(float), cudaMemcpyHostToDevice, e->stream));
    e->h_obs_rows = obs; e->h_slot_rows = nullptr;
    rc = search_impl(e, e->d_obs.as<float>(), stride, legal_mask, to_play, add_noise, n_trees, out);
    e->h_obs_rows = nullptr;
    return rc;
}

int mzb_search_batch_slots(mzb_engine* e, const float* obs, const int32_t* slots, const uint32_t* legal_mask,
                           const int32_t* to_play, const uint8_t* add_noise, int n_trees, mzb_search_result* out) {
    int rc = check_search_args(e, obs, legal_mask, to_play, add_noise, n_trees);
    if (rc) return rc;
    MZB_REQUIRE(slots, "null slot list");
    MZB_CUDA(cudaSetDevice(e->device));
    const int stride = (e->cfg.evaluator == MZB_EVAL_SYNTHETIC) ? 1 : e->dims.initial_flat();
    MZB_CUDA(cudaMemcpyAsync(e->d_obs.p, obs, (size_t)n_trees * stride * sizeof(float), cudaMemcpyHostToDevice, e->stream));
    e->h_obs_rows = obs; e->h_slot_rows = slots;
    rc = search_with_device_obs_slots(e, slots, legal_mask, to_play, add_noise, n_trees, out);
    e->h_obs_rows = nullptr; e->h_slot_rows = nullptr;
    return rc;
}

int mzb_search_batch_device_obs(mzb_engine* e, const float* d_obs, const uint32_t* legal_mask, const int32_t* to_play,
                                const uint8_t* add_noise, int n_trees, mzb_search_result* out) {
    int rc = check_search_args(e, d_obs, legal_mask, to_play, add_noise, n_trees);
    if (rc) return rc;
    MZB_CUDA(cudaSetDevice(e->device));
    const int stride = (e->cfg.evaluator == MZB_EVAL_SYNTHETIC) ? 1 : e->dims.initial_flat();
    return search_impl(e, d_obs, stride, legal_mask, to_play, add_noise, n_trees, out);
}

int mzb_last_search_timing(mzb_engine* e, mzb_search_timing* out) {
    MZB_REQUIRE(e && out, "null argument");
    *out = e->timing;
    return MZB_OK;
}

int mzb_search_get_trace(mzb_engine* e, int T, int32_t* parent, int32_t* action) {
    MZB_REQUIRE(e && parent && action, "null argument");
    MZB_REQUIRE(e->recording && e->trace_parent.p, "recording is not enabled (mzb_search_set_recording)");
    MZB_REQUIRE(T >= 1 && T <= e->cfg.max_trees, "bad n_trees");
    MZB_CUDA(cudaSetDevice(e->device));
    const size_t n = e->cfg.num_simulations;
    MZB_CUDA(cudaMemcpy(parent, e->trace_parent.p, T * n * sizeof(int), cudaMemcpyDeviceToHost));
    MZB_CUDA(cudaMemcpy(action, e->trace_action.p, T * n * sizeof(int), cudaMemcpyDeviceToHost));
    return MZB_OK;
}

int mzb_search_get_records(mzb_engine* e, int T, double* value, double* reward, double* policy) {
    MZB_REQUIRE(e && value && reward && policy, "null argument");
    MZB_REQUIRE(e->recording && e->rec_value.p, "recording is not enabled (mzb_search_set_recording)");
    MZB_REQUIRE(T >= 1 && T <= e->cfg.max_trees, "bad n_trees");
    MZB_CUDA(cudaSetDevice(e->device));
    const size_t n1 = e->cfg.num_simulations + 1, A = e->cfg.num_actions;
    MZB_CUDA(cudaMemcpy(value, e->rec_value.p, T * n1 * sizeof(double), cudaMemcpyDeviceToHost));
    MZB_CUDA(cudaMemcpy(reward, e->rec_reward.p, T * n1 * sizeof(double), cudaMemcpyDeviceToHost));
    MZB_CUDA(cudaMemcpy(policy, e->rec_policy.p, T * n1 * A * sizeof(double), cudaMemcpyDeviceToHost));
    return MZB_OK;
}

// ------------------------------------------------------------------ VPRNetModel batch API (host pointers)
static int infer_common(mzb_engine* e, int batch) {
    MZB_REQUIRE(e, "null argument");
    MZB_REQUIRE(e->net && e->weights_loaded, "weights not loaded (mzb_load_weights)");
    MZB_REQUIRE(batch >= 1, "empty batch");
    MZB_CUDA(cudaSetDevice(e->device));
    return MZB_OK;
}

int mzb_initial_inference(mzb_engine* e, const float* obs, int batch, double* value, double* reward, double* policy, float* hidden) {
    int rc = infer_common(e, batch);
    if (rc) return rc;
    MZB_REQUIRE(obs, "null observations");
    const int T = e->cfg.max_trees, A = e->cfg.num_actions, flat = e->dims.initial_flat(), S = e->dims.hidden_size();
    const int scratch0 = T * (e->cfg.num_simulations + 1);   // compat-path pool slots live behind the trees' slots
    cudaStream_t s = e->stream;
    if (!e->c_hidden_out.p) MZB_CUDA(e->c_hidden_out.alloc((size_t)T * S * sizeof(float)));
    for (int b0 = 0; b0 < batch; b0 += T) {
        const int B = (batch - b0 < T) ? batch - b0 : T;
        MZB_CUDA(cudaMemcpyAsync(e->d_obs.p, obs + (size_t)b0 * flat, (size_t)B * flat * sizeof(float), cudaMemcpyHostToDevice, s));
        const int* out_slots = e->d_iota.as<int>();   // iota values 0..2T-1; offset by scratch0 through a shifted copy below
        (void)out_slots;
        // slots scratch0 + [0, B): reuse d_in_slot storage as the index array
        std::vector<int> idx(B);
        for (int i = 0; i < B; ++i) idx[i] = scratch0 + i;
        MZB_CUDA(cudaMemcpyAsync(e->in_slot.p, idx.data(), B * sizeof(int), cudaMemcpyHostToDevice, s));
        MZB_CUDA(cudaStreamSynchronize(s));
        rc = e->net->initial(e->d_obs.as<float>(), B, e->in_slot.as<int>(), tree_outputs(e), s);
        if (rc) return rc;
        if (hidden) {
            rc = e->net->export_hidden(e->c_hidden_out.as<float>(), B, e->in_slot.as<int>(), s);
            if (rc) return rc;
            MZB_CUDA(cudaMemcpyAsync(hidden + (size_t)b0 * S, e->c_hidden_out.p, (size_t)B * S * sizeof(float), cudaMemcpyDeviceToHost, s));
        }
        if (value) MZB_CUDA(cudaMemcpyAsync(value + b0, e->td.value, B * sizeof(double), cudaMemcpyDeviceToHost, s));
        if (reward) MZB_CUDA(cudaMemcpyAsync(reward + b0, e->td.reward_out, B * sizeof(double), cudaMemcpyDeviceToHost, s));
        if (policy) MZB_CUDA(cudaMemcpyAsync(policy + (size_t)b0 * A, e->td.policy, (size_t)B * A * sizeof(double), cudaMemcpyDeviceToHost, s));
        MZB_CUDA(cudaStreamSynchronize(s));
    }
    return MZB_OK;
}

int mzb_recurrent_inference(mzb_engine* e, const float* hidden, const int32_t* action, int batch, double* value,
                            double* reward, double* policy, float* hidden_out) {
    int rc = infer_common(e, batch);
    if (rc) return rc;
    MZB_REQUIRE(hidden && action, "null inputs");
    const int T = e->cfg.max_trees, A = e->cfg.num_actions, S = e->dims.hidden_size();
    for (int i = 0; i < batch; ++i) MZB_REQUIRE(action[i] >= 0 && action[i] < A, "action %d of sample %d outside the action space", action[i], i);
    const int scratch0 = T * (e->cfg.num_simulations + 1);
    cudaStream_t s = e->stream;
    if (!e->c_hidden_in.p) MZB_CUDA(e->c_hidden_in.alloc((size_t)T * S * sizeof(float)));
    if (!e->c_hidden_out.p) MZB_CUDA(e->c_hidden_out.alloc((size_t)T * S * sizeof(float)));
    for (int b0 = 0; b0 < batch; b0 += T) {
        const int B = (batch - b0 < T) ? batch - b0 : T;
        std::vector<int> idx(2 * B);
        for (int i = 0; i < B; ++i) { idx[i] = scratch0 + i; idx[B + i] = scratch0 + T + i; }
        MZB_CUDA(cudaMemcpyAsync(e->in_slot.p, idx.data(), B * sizeof(int), cudaMemcpyHostToDevice, s));
        MZB_CUDA(cudaMemcpyAsync(e->out_slot.p, idx.data() + B, B * sizeof(int), cudaMemcpyHostToDevice, s));
        MZB_CUDA(cudaMemcpyAsync(e->action.p, action + b0, B * sizeof(int), cudaMemcpyHostToDevice, s));
        MZB_CUDA(cudaMemcpyAsync(e->c_hidden_in.p, hidden + (size_t)b0 * S, (size_t)B * S * sizeof(float), cudaMemcpyHostToDevice, s));
        MZB_CUDA(cudaStreamSynchronize(s));
        rc = e->net->import_hidden(e->c_hidden_in.as<float>(), B, e->in_slot.as<int>(), s);
        if (rc) return rc;
        rc = e->net->recurrent(B, e->in_slot.as<int>(), e->action.as<int>(), e->out_slot.as<int>(), tree_outputs(e), s);
        if (rc) return rc;
        if (hidden_out) {
            rc = e->net->export_hidden(e->c_hidden_out.as<float>(), B, e->out_slot.as<int>(), s);
            if (rc) return rc;
            MZB_CUDA(cudaMemcpyAsync(hidden_out + (size_t)b0 * S, e->c_hidden_out.p, (size_t)B * S * sizeof(float), cudaMemcpyDeviceToHost, s));
        }
        if (value) MZB_CUDA(cudaMemcpyAsync(value + b0, e->td.value, B * sizeof(double), cudaMemcpyDeviceToHost, s));
        if (reward) MZB_CUDA(cudaMemcpyAsync(reward + b0, e->td.reward_out, B * sizeof(double), cudaMemcpyDeviceToHost, s));
        if (policy) MZB_CUDA(cudaMemcpyAsync(policy + (size_t)b0 * A, e->td.policy, (size_t)B * A * sizeof(double), cudaMemcpyDeviceToHost, s));
        MZB_CUDA(cudaStreamSynchronize(s));
    }
    return MZB_OK;
}

int mzb_bench_recurrent(mzb_engine* e, int batch, int iters, float* ms_per_pass, int* launches_per_pass) {
    int rc = infer_common(e, batch);
    if (rc) return rc;
    const int T = e->cfg.max_trees, A = e->cfg.num_actions, S = e->dims.hidden_size(), n = e->cfg.num_simulations;
    MZB_REQUIRE(batch <= T && iters >= 1, "batch above max_trees or iters < 1");
    cudaStream_t s = e->stream;
    // device-resident inputs: pseudo-random hidden states in [0,1) and actions, imported once into pool block 0.  Pass i reads
    // block i % (n + 1) and writes block (i + 1) % (n + 1) of the trees' node pool (blocks of `batch` slots), exactly like
    // simulation i of a search: every pass consumes the previous pass's states, and the working set is the whole pool
    // ((n + 1) x batch states, 0.58 GB for C4d at 1024), larger than the 126 MB L2 — no L2-resident re-reads between passes.
    if (!e->c_hidden_in.p) MZB_CUDA(e->c_hidden_in.alloc((size_t)T * S * sizeof(float)));
    std::vector<float> h((size_t)batch * S);
    uint32_t x = 12345u;
    for (auto& v : h) { x = x * 1664525u + 1013904223u; v = (float)(x >> 8) / 16777216.0f; }
    const int nb = n + 1;
    std::vector<int> idx((size_t)nb * batch), act(batch);
    for (size_t i = 0; i < idx.size(); ++i) idx[i] = (int)i;
    for (int i = 0; i < batch; ++i) act[i] = i % A;
    if (e->bench_slots.bytes < idx.size() * sizeof(int)) MZB_CUDA(e->bench_slots.alloc(idx.size() * sizeof(int)));
    MZB_CUDA(cudaMemcpy(e->c_hidden_in.p, h.data(), h.size() * sizeof(float), cudaMemcpyHostToDevice));
    MZB_CUDA(cudaMemcpy(e->bench_slots.p, idx.data(), idx.size() * sizeof(int), cudaMemcpyHostToDevice));
    MZB_CUDA(cudaMemcpy(e->action.p, act.data(), batch * sizeof(int), cudaMemcpyHostToDevice));
    const int* slots = e->bench_slots.as<int>();
    rc = e->net->import_hidden(e->c_hidden_in.as<float>(), batch, slots, s);
    if (rc) return rc;
    int blk = 0;
    auto pass = [&]() -> int {
        const int nxt = (blk + 1) % nb;
        int r = e->net->recurrent(batch, slots + (size_t)blk * batch, e->action.as<int>(), slots + (size_t)nxt * batch, tree_outputs(e), s);
        blk = nxt;
        return r;
    };
    for (int w = 0; w < 3; ++w) { rc = pass(); if (rc) return rc; }
    MZB_CUDA(cudaEventRecord(e->ev_start, s));
    for (int i = 0; i < iters; ++i) { rc = pass(); if (rc) return rc; }
    MZB_CUDA(cudaEventRecord(e->ev_end, s));
    MZB_CUDA(cudaStreamSynchronize(s));
    float ms = 0.f;
    MZB_CUDA(cudaEventElapsedTime(&ms, e->ev_start, e->ev_end));
    if (ms_per_pass) *ms_per_pass = ms / iters;
    if (launches_per_pass) *launches_per_pass = e->net->launches_recurrent();
    return MZB_OK;
}

}  // extern "C"
