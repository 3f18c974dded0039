// tree.cu — warp-per-tree MCTS kernels (select / expand / backup / root set-up / statistics) and the SE stub.
// Compile with -fmad=false: every f64 expression below must round exactly like the reference's x86-64 SSE2 code
// (muzero-cpp/mcts.cpp:90-141, built -O3 without -march => no FMA).
//
// HBM layout: [tree][slot] structure-of-arrays; a node's children occupy consecutive slots, so a warp loads the
// child statistics of one node with one coalesced request per array (lane j <-> child j) and the pUCT arg-max is a
// warp-shuffle reduction.  Integer RNG (std::mt19937 + libstdc++'s std::shuffle / Lemire downscaling) is restated
// on the device so that the per-visit child shuffle (mcts.cpp:112) is reproduced bit for bit.
#include "tree.h"
#include "umma.cuh"

namespace mzb {
namespace {

constexpr unsigned FULL = 0xffffffffu;

// ------------------------------------------------------------------ std::mt19937 for one warp
struct WarpRng {
    volatile uint32_t* mt;
    int idx;          // position of the next unconsumed state word (624 => twist first), like libstdc++'s _M_p
    uint32_t window;  // lane l holds the tempered output for position wbase + l
    int wcount, wpos;
    int lane;

    __device__ void init(uint32_t* state, int position) {
        mt = state; idx = position; wcount = 0; wpos = 0; lane = threadIdx.x & 31;
    }
    __device__ static uint32_t twist_word(uint32_t cur, uint32_t nxt, uint32_t far) {
        uint32_t y = (cur & 0x80000000u) | (nxt & 0x7fffffffu);
        return far ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
    }
    // mersenne_twister_engine::_M_gen_rand, 32 words per step; reads complete before the step's writes
    __device__ void twist() {
        for (int base = 0; base < 227; base += 32) {
            int k = base + lane;
            uint32_t v = 0;
            bool on = k < 227;
            if (on) v = twist_word(mt[k], mt[k + 1], mt[k + 397]);
            __syncwarp();
            if (on) mt[k] = v;
            __syncwarp();
        }
        for (int base = 227; base < 623; base += 32) {
            int k = base + lane;
            uint32_t v = 0;
            bool on = k < 623;
            if (on) v = twist_word(mt[k], mt[k + 1], mt[k - 227]);
            __syncwarp();
            if (on) mt[k] = v;
            __syncwarp();
        }
        if (lane == 0) mt[623] = twist_word(mt[623], mt[0], mt[396]);
        __syncwarp();
    }
    __device__ void refill() {
        if (idx >= 624) { twist(); idx = 0; }
        wcount = 624 - idx; if (wcount > 32) wcount = 32;
        wpos = 0;
        uint32_t y = (lane < wcount) ? mt[idx + lane] : 0u;
        y ^= y >> 11;
        y ^= (y << 7) & 0x9d2c5680u;
        y ^= (y << 15) & 0xefc60000u;
        y ^= y >> 18;
        window = y;
    }
    __device__ uint32_t next() {  // warp-uniform
        if (wpos >= wcount) refill();
        uint32_t v = __shfl_sync(FULL, window, wpos);
        ++wpos; ++idx;
        return v;
    }
    // uniform_int_distribution<unsigned long>{0, range-1}: Lemire's method on 32-bit draws (uniform_int_dist.h:252-281)
    __device__ uint32_t below(uint32_t range) {
        unsigned long long product = (unsigned long long)next() * (unsigned long long)range;
        uint32_t low = (uint32_t)product;
        if (low < range) {
            uint32_t threshold = (0u - range) % range;
            while (low < threshold) {
                product = (unsigned long long)next() * (unsigned long long)range;
                low = (uint32_t)product;
            }
        }
        return (uint32_t)(product >> 32);
    }
};

// ceil(65536 / b) for b = 0 .. 33: x / b == (x * kRecip16[b]) >> 16 exactly for x * (b - 1) < 65536 — std::shuffle's paired
// draw is x < (i + 1)(i + 2) <= 1056 with b = i + 2 <= 33.  Replaces two runtime integer divisions (~240 warp instructions per
// tree and step, the largest single item in the ncu source view).
__constant__ uint32_t kRecip16[34] = {0, 65536, 32768, 21846, 16384, 13108, 10923, 9363, 8192, 7282, 6554, 5958, 5462, 5042, 4682, 4370,
                                      4096, 3856, 3641, 3450, 3277, 3121, 2979, 2850, 2731, 2622, 2521, 2428, 2341, 2260, 2185, 2115, 2048, 1986};
// std::shuffle (stl_algo.h:3741-3800) on a permutation distributed over lanes: lane j holds perm[j]
__device__ __forceinline__ int swap_lanes(int perm, int a, int b, int lane) {
    int va = __shfl_sync(FULL, perm, a);
    int vb = __shfl_sync(FULL, perm, b);
    if (lane == a) perm = vb; else if (lane == b) perm = va;
    return perm;
}
__device__ int shuffle_perm(WarpRng& rng, int n, int lane) {
    int perm = lane;
    if (n <= 1) return perm;
    int i = 1;
    if ((n & 1) == 0) {
        int j = (int)rng.below(2u);
        perm = swap_lanes(perm, i, j, lane);
        ++i;
    }
    while (i != n) {
        uint32_t swap_range = (uint32_t)i + 1u;
        uint32_t b1 = swap_range + 1u;
        uint32_t x = rng.below(swap_range * b1);
        int p0, p1;
        if (b1 <= 33u) { p0 = (int)((x * kRecip16[b1]) >> 16); p1 = (int)(x - (uint32_t)p0 * b1); }
        else { p0 = (int)(x / b1); p1 = (int)(x % b1); }
        perm = swap_lanes(perm, i, p0, lane); ++i;
        perm = swap_lanes(perm, i, p1, lane); ++i;
    }
    return perm;
}

__device__ __forceinline__ double shfl_d(double v, int src) { return __shfl_sync(FULL, v, src); }

// ------------------------------------------------------------------ root set-up (mcts.cpp:39-49)
__global__ void tree_root_kernel(TreeDev td, int n_trees, const uint32_t* legal_mask, const int* to_play,
                                 const uint8_t* add_noise, const double* noise) {
    const int t = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (t >= n_trees) return;
    const long sb = (long)t * td.slots_per_tree;
    const uint32_t mask = legal_mask[t];
    const int L = __popc(mask);
    // root node: MCTSNode root(InvalidAction, 0); expand(legal, to_play, reward, policy)
    if (lane == 0) {
        td.meta[sb] = make_int4(0, -1, 1, 0);
        td.prior[sb] = 0.0; td.value_sum[sb] = 0.0; td.reward[sb] = td.reward_out[t];
        td.mm_min[t] = td.init_min; td.mm_max[t] = td.init_max;   // minmax_stats_.reset()
        td.root_to_play[t] = to_play[t]; td.n_legal[t] = L; td.n_slots[t] = 1 + L; td.max_depth[t] = 0;
        if (td.rec_value) {
            td.rec_value[(long)t * (td.n_sims + 1)] = td.value[t];
            td.rec_reward[(long)t * (td.n_sims + 1)] = td.reward_out[t];
        }
    }
    if (td.rec_policy && lane < td.A) td.rec_policy[(long)t * (td.n_sims + 1) * td.A + lane] = td.policy[(long)t * td.A + lane];
    // lane j <- j-th legal action in ascending order (children in legal_actions order, mcts.cpp:163-166)
    if (lane < L) {
        uint32_t m = mask;
        for (int k = 0; k < lane; ++k) m &= m - 1;
        const int a = __ffs(m) - 1;
        double p = td.policy[(long)t * td.A + a];
        if (add_noise[t]) {  // MCTSNode::add_exploration_noise (mcts.cpp:170-176)
            const double eps = td.dirichlet_epsilon;
            const double keep = 1 - eps;
            const double left = p * keep;
            const double right = noise[(long)t * td.A + lane] * eps;
            p = left + right;
        }
        td.meta[sb + 1 + lane] = make_int4(0, a, -1, -1);
        td.prior[sb + 1 + lane] = p;
        td.value_sum[sb + 1 + lane] = 0.0;
        td.reward[sb + 1 + lane] = 0.0;
    }
}

// ------------------------------------------------------------------ one simulation step
__global__ void __launch_bounds__(128) tree_step_kernel(TreeDev td, int t_first, int t_end, int sim) {
    const int t = t_first + ((blockIdx.x * blockDim.x + threadIdx.x) >> 5);
    const int lane = threadIdx.x & 31;
    if (t >= t_end) return;
    const long sb = (long)t * td.slots_per_tree;
    int4* meta = td.meta + sb;
    double* prior = td.prior + sb;
    double* value_sum = td.value_sum + sb;
    double* reward = td.reward + sb;
    int* path = td.path_slot + (long)t * (td.n_sims + 2);
    const int players = td.players;
    const int root_tp = td.root_to_play[t];
    const double discount = td.discount;
    double mm_min = td.mm_min[t], mm_max = td.mm_max[t];

    if (sim > 0) {
        // ---- expand the leaf chosen by the previous step with the network outputs (mcts.cpp:72-75) ----
        const int len = td.path_len[t];
        const int leaf = path[len - 1];
        const int vtp = (root_tp + (len - 1)) % players;  // virtual_to_play after len-1 descents (mcts.cpp:65)
        const int base = td.n_slots[t];
        const double net_value = td.value[t];
        const double net_reward = td.reward_out[t];
        if (lane < td.A) {
            meta[base + lane] = make_int4(0, lane, -1, -1);
            prior[base + lane] = td.policy[(long)t * td.A + lane];
            value_sum[base + lane] = 0.0;
            reward[base + lane] = 0.0;
        }
        if (td.rec_policy && lane < td.A)
            td.rec_policy[((long)t * (td.n_sims + 1) + sim) * td.A + lane] = td.policy[(long)t * td.A + lane];
        if (lane == 0) {
            td.n_slots[t] = base + td.A;
            if (td.rec_value) {
                td.rec_value[(long)t * (td.n_sims + 1) + sim] = net_value;
                td.rec_reward[(long)t * (td.n_sims + 1) + sim] = net_reward;
            }
        }
        // ---- backup along the search path, leaf -> root (mcts.cpp:90-107) ----
        double value = net_value;
        for (int chunk = (len - 1) / 32; chunk >= 0; --chunk) {
            const int k = chunk * 32 + lane;
            int slot = 0; int4 m = make_int4(0, 0, 0, 0); double vs = 0.0, r = 0.0;
            if (k < len) {
                slot = path[k];
                m = meta[slot]; vs = value_sum[slot];
                r = (k == len - 1) ? net_reward : reward[slot];
            }
            const int hi = min(31, len - 1 - chunk * 32);
            for (int l = hi; l >= 0; --l) {
                const int kk = chunk * 32 + l;
                const double vs_l = shfl_d(vs, l);
                const double r_l = shfl_d(r, l);
                const int vc_l = __shfl_sync(FULL, m.x, l);
                const int tp_k = players == 2 ? ((root_tp + kk) & 1) : (players == 1 ? 0 : (root_tp + kk) % players);
                const bool same = tp_k == vtp;
                const double nvs = vs_l + (same ? value : -value);
                const int nvc = vc_l + 1;
                const double nodeval = nvs / nvc;
                const double signedv = (players == 2) ? -nodeval : nodeval;
                const double dq = discount * signedv;
                const double new_q = r_l + dq;
                mm_max = (mm_max < new_q) ? new_q : mm_max;
                mm_min = (new_q < mm_min) ? new_q : mm_min;
                const double rr = (same && players == 2) ? -r_l : r_l;
                const double dv = discount * value;
                value = rr + dv;
                if (lane == l) { vs = nvs; m.x = nvc; }
            }
            if (k < len) {
                if (k == len - 1) { m.z = base; m.w = sim; reward[slot] = net_reward; }
                meta[slot] = m;
                value_sum[slot] = vs;
            }
        }
        if (lane == 0) { td.mm_min[t] = mm_min; td.mm_max[t] = mm_max; }
        __syncwarp();
    }
    if (sim >= td.n_sims) return;

    // ---- select: descend with pUCT until an unexpanded node (mcts.cpp:60-66,110-141) ----
    WarpRng rng;
    rng.init(td.mt + (long)t * 624, td.mt_idx[t]);
    int4 cur = meta[0];
    int slot = 0, len = 1, n = td.n_legal[t];
    int parent_node = 0;
    if (lane == 0) path[0] = 0;
    const bool normalize = mm_max > mm_min;
    const double mm_range = mm_max - mm_min;
    while (cur.z >= 0) {
        const int cb = cur.z;
        const int N = cur.x;
        parent_node = cur.w;
        // coalesced child-statistics loads: lane j <-> child j
        int4 cm = make_int4(0, 0, -1, -1); double cp = 0.0, cvs = 0.0, cr = 0.0;
        if (lane < n) { cm = meta[cb + lane]; cp = prior[cb + lane]; cvs = value_sum[cb + lane]; cr = reward[cb + lane]; }
        const double pbc_n = __ldg(td.pbc_table + N);   // in flight with the child loads, consumed after the shuffle
        // sqrt(parent visits): correctly rounded on both sides, so the host-built table is bit-identical to the software fp64
        // sqrt it replaces (hand-built trees of the unit-test hooks may exceed the table)
        const double sqrt_n = (N <= td.n_sims + 1) ? __ldg(td.pbc_table + (td.n_sims + 2) + N) : sqrt((double)N);
        // std::shuffle(children) — physically permutes the child records
        const int perm = shuffle_perm(rng, n, lane);
        int4 pm;
        pm.x = __shfl_sync(FULL, cm.x, perm); pm.y = __shfl_sync(FULL, cm.y, perm);
        pm.z = __shfl_sync(FULL, cm.z, perm); pm.w = __shfl_sync(FULL, cm.w, perm);
        const double pp = shfl_d(cp, perm), pvs = shfl_d(cvs, perm), pr = shfl_d(cr, perm);
        if (lane < n) { meta[cb + lane] = pm; prior[cb + lane] = pp; value_sum[cb + lane] = pvs; reward[cb + lane] = pr; }
        // ucb_score (mcts.cpp:127-141)
        double score = 0.0;
        if (lane < n) {
            double pb_c = pbc_n;
            const double ratio = sqrt_n / (double)(pm.x + 1);
            pb_c = pb_c * ratio;
            const double prior_score = pb_c * pp;
            double value_score = 0.0;
            if (pm.x > 0) {
                const double cval = pvs / pm.x;
                const double child_value = (players == 1) ? cval : -cval;
                const double dq = discount * child_value;
                const double q = pr + dq;
                value_score = normalize ? (q - mm_min) / mm_range : q;
            }
            score = prior_score + value_score;
        }
        if (td.dbg_score && len == 1) td.dbg_score[lane] = score;
        // first strict arg-max over children 0..n-1 (mcts.cpp:115-121)
        // (an index >= 64 marks a lane without a child; the butterfly only spans the lanes that can hold one)
        int best = lane < n ? lane : 64 + lane;
        double bs = score;
        const int off0 = n <= 8 ? 4 : (n <= 16 ? 8 : 16);   // warp-uniform
        for (int off = off0; off > 0; off >>= 1) {
            const double os = __shfl_xor_sync(FULL, bs, off);
            const int oi = __shfl_xor_sync(FULL, best, off);
            if (oi < 64 && (best >= 64 || os > bs || (os == bs && oi < best))) { bs = os; best = oi; }
        }
        best = __shfl_sync(FULL, best, 0);   // lanes beyond the butterfly's span take lane 0's result
        cur.x = __shfl_sync(FULL, pm.x, best); cur.y = __shfl_sync(FULL, pm.y, best);
        cur.z = __shfl_sync(FULL, pm.z, best); cur.w = __shfl_sync(FULL, pm.w, best);
        slot = cb + best;
        if (lane == 0) path[len] = slot;
        ++len;
        n = td.A;
    }
    if (lane == 0) {
        td.mt_idx[t] = rng.idx;
        td.path_len[t] = len;
        const int depth = len - 1;
        if (depth > td.max_depth[t]) td.max_depth[t] = depth;
        const int pool_base = t * (td.n_sims + 1);
        td.in_slot[t] = pool_base + parent_node;
        td.out_slot[t] = pool_base + sim + 1;
        td.action[t] = cur.y;
        if (td.trace_parent) {
            td.trace_parent[(long)t * td.n_sims + sim] = parent_node;
            td.trace_action[(long)t * td.n_sims + sim] = cur.y;
        }
    }
}

// ------------------------------------------------------------------ get_search_statistics (mcts.cpp:179-200)
__global__ void tree_stats_kernel(TreeDev td, int n_trees, ResultDev r) {
    const int t = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (t >= n_trees) return;
    const long sb = (long)t * td.slots_per_tree;
    const int L = td.n_legal[t];
    int4 cm = make_int4(0, -1, -1, -1);
    if (lane < L) cm = td.meta[sb + 1 + lane];
    int sum = (lane < L) ? cm.x : 0;
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) sum += __shfl_xor_sync(FULL, sum, off);
    const long ob = (long)t * td.A;
    if (lane < td.A) {
        r.child_actions[ob + lane] = (lane < L) ? cm.y : -1;
        r.child_visits[ob + lane] = (lane < L) ? cm.x : 0;
        r.relative_visit[ob + lane] = (lane < L) ? (double)cm.x / sum : 0.0;
        r.emperical_policy[ob + lane] = 0.0 / sum;  // actions without a child: 0 / sum_visits
    }
    __syncwarp();
    if (lane < L) r.emperical_policy[ob + cm.y] = (double)cm.x / sum;
    if (lane == 0) {
        const int4 root = td.meta[sb];
        r.root_value[t] = (root.x == 0) ? 0.0 : td.value_sum[sb] / root.x;
        r.max_tree_depth[t] = td.max_depth[t];
        r.n_children[t] = L;
    }
}

// ------------------------------------------------------------------ SE synthetic evaluator
__device__ __forceinline__ uint32_t se_mix(uint32_t x) {
    x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16;
    return x;
}
// Whole warp: lane i produces policy weight i.  The weights are integers in [1, 1024] held in f64, so their sum is exact in any
// order and equals a sequential sum bit for bit.
__device__ void se_outputs(const TreeDev& td, const SEDev& se, int t, uint32_t cid, int lane) {
    if (lane == 0) {
        const double u = (double)(se_mix(cid ^ 0x1111U) >> 8) / 16777216.0;
        const double vspan = se.vhi - se.vlo;
        const double vs = u * vspan;
        td.value[t] = se.vlo + vs;
        const double u2 = (double)(se_mix(cid ^ 0x2222U) >> 8) / 16777216.0;
        const double rspan = se.rhi - se.rlo;
        const double rs = u2 * rspan;
        td.reward_out[t] = se.rlo + rs;
    }
    const double w = (lane < td.A) ? (double)(1U + (se_mix(cid ^ (0x3333U + 0x101U * (uint32_t)lane)) & 0x3FFU)) : 0.0;
    double sum = w;
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) sum += __shfl_xor_sync(FULL, sum, off);
    if (lane < td.A) td.policy[(long)t * td.A + lane] = w / sum;
}
__global__ void se_initial_kernel(TreeDev td, SEDev se, int n_trees, const float* obs, int obs_stride) {
    const int t = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (t >= n_trees) return;
    uint32_t key;
    if (obs_stride == 1) {
        key = (uint32_t)obs[t];
    } else {   // MZB_EVAL_SYNTHETIC_OBS: hash of the whole row — an order-independent sum of per-position hashes of the float bit patterns
        const float* row = obs + (long)t * obs_stride;
        key = 0;
        for (int i = lane; i < obs_stride; i += 32) key += se_mix(__float_as_uint(row[i]) + 0x9e3779b9U * (uint32_t)(i + 1));
        for (int off = 16; off > 0; off >>= 1) key += __shfl_xor_sync(FULL, key, off);
    }
    const uint32_t cid = se_mix(key ^ 0x9e3779b9U) & 0xFFFFFFU;
    if (lane == 0) se.node_cid[(long)t * (td.n_sims + 1)] = cid;
    se_outputs(td, se, t, cid, lane);
    float* dst = se.pool + (long)t * (td.n_sims + 1) * se.slot_floats;
    for (int i = lane; i < se.slot_floats; i += 32) dst[i] = (float)((cid + i) & 0xFF) * (1.0f / 256.0f);
}
__global__ void se_recurrent_kernel(TreeDev td, SEDev se, int n_trees) {
    const int t = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (t >= n_trees) return;
    const int in_slot = td.in_slot[t], out_slot = td.out_slot[t];
    const int pool_base = t * (td.n_sims + 1);
    const uint32_t pc = se.node_cid[in_slot];
    const uint32_t cid = se_mix(pc * 31U + (uint32_t)td.action[t] + 0x85ebca6bU) & 0xFFFFFFU;
    (void)pool_base;
    if (lane == 0) se.node_cid[out_slot] = cid;
    se_outputs(td, se, t, cid, lane);
    if (se.bulk) {   // the hidden state travels by bulk copy (se_move_kernel, on its own stream)
        if (lane == 0) { se.mv_in[t] = in_slot; se.mv_out[t] = out_slot; }
        return;
    }
    // odd state sizes: 16 B per lane per request (S_h read + S_h write)
    const float4* src = reinterpret_cast<const float4*>(se.pool + (long)in_slot * se.slot_floats);
    float4* dst = reinterpret_cast<float4*>(se.pool + (long)out_slot * se.slot_floats);
    const int n4 = (se.slot_floats % 4 == 0) ? (se.slot_floats >> 2) : 0;
    for (int i = lane; i < n4; i += 32) dst[i] = src[i];
    for (int i = (n4 << 2) + lane; i < se.slot_floats; i += 32) se.pool[(long)out_slot * se.slot_floats + i] = se.pool[(long)in_slot * se.slot_floats + i];
}

// The stub's hidden-state move, parent slot -> child slot, staged through shared memory: one CTA per tree, one elected thread,
// two 1-D bulk copies (global -> smem on an mbarrier, smem -> global as a bulk group).  The copy engine moves the 16-byte
// lines; no register traffic.  ~10 CTAs of a Connect4 state (21.5 KB) are resident per SM, so 1024+ trees are all in flight.
__global__ void __launch_bounds__(32) se_move_kernel(SEDev se, int n_trees) {
    extern __shared__ __align__(128) unsigned char stage[];
    __shared__ __align__(8) uint64_t bar;
    const int t = blockIdx.x;
    if (t >= n_trees || threadIdx.x != 0) return;
    const uint32_t bytes = (uint32_t)se.slot_floats * 4u;
    const float* src = se.pool + (long)se.mv_in[t] * se.slot_floats;
    float* dst = se.pool + (long)se.mv_out[t] * se.slot_floats;
    umma::mbar_init(&bar, 1);
    umma::fence_mbar_init();
    umma::mbar_expect_tx(&bar, bytes);
    umma::bulk_g2s(stage, src, bytes, &bar);
    umma::mbar_wait(&bar, 0);
    umma::bulk_s2g(dst, stage, bytes);
    umma::bulk_commit_wait_all();
}

__global__ void mt_seed_kernel(uint32_t* mt, int* idx, const int* seeds, int first_slot, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t* s = mt + (long)(first_slot + i) * 624;
    uint32_t x = (uint32_t)seeds[i];
    s[0] = x;
    for (int k = 1; k < 624; ++k) { x = 1812433253U * (x ^ (x >> 30)) + (uint32_t)k; s[k] = x; }
    idx[first_slot + i] = 624;
}

__global__ void mt_move_kernel(uint32_t* home, int* idx_home, uint32_t* rows, int* idx_rows, const int* slots, int n, int dir) {
    const int i = blockIdx.x;
    if (i >= n) return;
    uint32_t* h = home + (long)slots[i] * 624;
    uint32_t* r = rows + (long)i * 624;
    for (int k = threadIdx.x; k < 624; k += blockDim.x) { if (dir == 0) r[k] = h[k]; else h[k] = r[k]; }
    if (threadIdx.x == 0) { if (dir == 0) idx_rows[i] = idx_home[slots[i]]; else idx_home[slots[i]] = idx_rows[i]; }
}

inline unsigned warps_grid(int n_trees, int block) { return (unsigned)(((long)n_trees * 32 + block - 1) / block); }

}  // namespace

void launch_tree_root(const TreeDev& td, int n_trees, const uint32_t* d_legal_mask, const int* d_to_play,
                      const uint8_t* d_add_noise, const double* d_noise, cudaStream_t s) {
    tree_root_kernel<<<warps_grid(n_trees, 128), 128, 0, s>>>(td, n_trees, d_legal_mask, d_to_play, d_add_noise, d_noise);
}
void launch_tree_step(const TreeDev& td, int n_trees, int sim, cudaStream_t s) {
    tree_step_kernel<<<warps_grid(n_trees, 128), 128, 0, s>>>(td, 0, n_trees, sim);
}
void launch_tree_step_range(const TreeDev& td, int t_first, int t_end, int sim, cudaStream_t s) {
    tree_step_kernel<<<warps_grid(t_end - t_first, 128), 128, 0, s>>>(td, t_first, t_end, sim);
}
void launch_tree_stats(const TreeDev& td, int n_trees, const ResultDev& r, cudaStream_t s) {
    tree_stats_kernel<<<warps_grid(n_trees, 128), 128, 0, s>>>(td, n_trees, r);
}
void launch_se_initial(const TreeDev& td, const SEDev& se, int n_trees, const float* d_obs, int obs_stride, cudaStream_t s) {
    se_initial_kernel<<<warps_grid(n_trees, 128), 128, 0, s>>>(td, se, n_trees, d_obs, obs_stride);
}
void launch_se_recurrent(const TreeDev& td, const SEDev& se, int n_trees, cudaStream_t s) {
    se_recurrent_kernel<<<warps_grid(n_trees, 128), 128, 0, s>>>(td, se, n_trees);
}
bool se_move_available(int slot_floats) { return slot_floats % 4 == 0 && (size_t)slot_floats * 4 <= 200 * 1024; }
void launch_se_move(const SEDev& se, int n_trees, cudaStream_t s) {
    const size_t bytes = (size_t)se.slot_floats * 4;
    // states above the 48 KB default need the opt-in (ALE: 18 KB, C4d: 21.5 KB do not).  The attribute is per DEVICE (a
    // DeviceManager may hold engines on several), so it is set on every such launch for the current one: a host-side store.
    if (bytes > 48 * 1024) cudaFuncSetAttribute(se_move_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    se_move_kernel<<<n_trees, 32, bytes, s>>>(se, n_trees);
}
void launch_mt_move(uint32_t* d_mt_home, int* d_idx_home, uint32_t* d_mt_rows, int* d_idx_rows, const int* d_slots, int n,
                    int dir, cudaStream_t s) {
    mt_move_kernel<<<n, 128, 0, s>>>(d_mt_home, d_idx_home, d_mt_rows, d_idx_rows, d_slots, n, dir);
}
void launch_mt_seed(uint32_t* d_mt, int* d_idx, const int* d_seeds, int first_slot, int n, cudaStream_t s) {
    mt_seed_kernel<<<(n + 127) / 128, 128, 0, s>>>(d_mt, d_idx, d_seeds, first_slot, n);
}

}  // namespace mzb
