/* se_evaluator.h — TEST INFRASTRUCTURE (oracle side).
 *
 * "SE", the synthetic evaluator used to pin search parity: a deterministic stand-in for
 * Evaluator::{InitialInference,RecurrentInference} (muzero-cpp/vprnet_evaluator.h:35-41), richer than the
 * reference's own fake (tests/mcts_test.cpp:16-27) so that random sweeps exercise ties, sign changes and
 * MinMax updates.  Every output is produced by integer hashing followed by IEEE-754 double operations
 * that round identically on x86-64 (no FMA: the oracle is built without -march) and on the GPU (tree
 * kernels are built with -fmad=false), so "identical network outputs" holds bit for bit on both sides.
 *
 *   content id (cid): 24-bit integer, exact in a float.
 *   root:   cid = mix(key ^ 0x9e3779b9) & 0xFFFFFF          key = (uint32) observation[0]
 *   child:  cid' = mix(cid * 31 + action + 0x85ebca6b) & 0xFFFFFF
 *   value  = lo + (mix(cid ^ 0x1111) >> 8) / 2^24 * (hi - lo)
 *   reward = lo + (mix(cid ^ 0x2222) >> 8) / 2^24 * (hi - lo)
 *   policy_i = w_i / sum_j w_j,  w_i = 1 + (mix(cid ^ (0x3333 + 0x101 * i)) & 0x3FF)   (sum in index order)
 */
#ifndef MZB_ORACLE_SE_EVALUATOR_H_
#define MZB_ORACLE_SE_EVALUATOR_H_
#include <stdint.h>

static inline uint32_t se_mix(uint32_t x) {
    x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16;
    return x;
}
/* key of a whole observation row (MZB_EVAL_SYNTHETIC_OBS): order-independent sum of per-position hashes of the float bit
 * patterns, so that a warp can form it with a shuffle reduction and the host with a plain loop */
static inline uint32_t se_obs_key(const float* obs, int n) {
    uint32_t key = 0;
    for (int i = 0; i < n; ++i) {
        union { float f; uint32_t u; } b;
        b.f = obs[i];
        key += se_mix(b.u + 0x9e3779b9U * (uint32_t)(i + 1));
    }
    return key;
}
static inline uint32_t se_initial_cid(uint32_t key) { return se_mix(key ^ 0x9e3779b9U) & 0xFFFFFFU; }
static inline uint32_t se_child_cid(uint32_t cid, uint32_t action) {
    return se_mix(cid * 31U + action + 0x85ebca6bU) & 0xFFFFFFU;
}
static inline double se_unit(uint32_t h) { return (double)(h >> 8) / 16777216.0; }
static inline double se_value(uint32_t cid, double lo, double hi) {
    double u = se_unit(se_mix(cid ^ 0x1111U));
    double span = hi - lo;
    double s = u * span;
    return lo + s;
}
static inline double se_reward(uint32_t cid, double lo, double hi) {
    double u = se_unit(se_mix(cid ^ 0x2222U));
    double span = hi - lo;
    double s = u * span;
    return lo + s;
}
static inline void se_policy(uint32_t cid, int num_actions, double* out) {
    double sum = 0.0;
    for (int i = 0; i < num_actions; ++i) {
        uint32_t w = 1U + (se_mix(cid ^ (0x3333U + 0x101U * (uint32_t)i)) & 0x3FFU);
        out[i] = (double)w;
        sum = sum + out[i];
    }
    for (int i = 0; i < num_actions; ++i) out[i] = out[i] / sum;
}
/* legal-action masks for root positions.  mode 0: all legal.  mode 1: pseudo-random non-empty subset. */
static inline uint32_t se_legal_mask(uint32_t key, int num_actions, int mode) {
    uint32_t full = (num_actions >= 32) ? 0xFFFFFFFFU : ((1U << num_actions) - 1U);
    if (mode == 0) return full;
    uint32_t m = se_mix(key ^ 0x5bd1e995U) & full;
    if (m == 0) m = 1U << (key % (uint32_t)num_actions);
    return m;
}
#endif
