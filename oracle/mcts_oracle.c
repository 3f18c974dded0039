/* mcts_oracle.c — TEST INFRASTRUCTURE.  See mcts_oracle.h.  Build: gcc -O2 -ffp-contract=off -fPIC -shared -lm.
 * Every function cites the reference (muzero-cpp/...) or libstdc++ 13.3 (bits/...) lines it restates. */
#include "mcts_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

#include "se_evaluator.h"

/* ------------------------------------------------------------------ std::mt19937 */
void mzo_mt_seed(mzo_mt19937* g, uint32_t seed) {
    g->mt[0] = seed;
    for (int i = 1; i < 624; ++i) g->mt[i] = 1812433253U * (g->mt[i - 1] ^ (g->mt[i - 1] >> 30)) + (uint32_t)i;
    g->idx = 624;
}
static void mt_twist(mzo_mt19937* g) {
    uint32_t* mt = g->mt;
    for (int k = 0; k < 624; ++k) {
        uint32_t y = (mt[k] & 0x80000000U) | (mt[(k + 1) % 624] & 0x7fffffffU);
        mt[k] = mt[(k + 397) % 624] ^ (y >> 1) ^ ((y & 1U) ? 0x9908b0dfU : 0U);
    }
    g->idx = 0;
}
uint32_t mzo_mt_next(mzo_mt19937* g) {
    if (g->idx >= 624) mt_twist(g);
    uint32_t y = g->mt[g->idx++];
    y ^= y >> 11;
    y ^= (y << 7) & 0x9d2c5680U;
    y ^= (y << 15) & 0xefc60000U;
    y ^= y >> 18;
    return y;
}

/* ------------------------------------------------------------------ uniform_int_distribution / shuffle */
/* bits/uniform_int_dist.h:252-281 (_S_nd<uint64_t>) reached through :323-329 because mt19937's range is 2^32-1 */
uint32_t mzo_uniform_below(mzo_mt19937* g, uint32_t range) {
    uint64_t product = (uint64_t)mzo_mt_next(g) * (uint64_t)range;
    uint32_t low = (uint32_t)product;
    if (low < range) {
        uint32_t threshold = (0U - range) % range;
        while (low < threshold) {
            product = (uint64_t)mzo_mt_next(g) * (uint64_t)range;
            low = (uint32_t)product;
        }
    }
    return (uint32_t)(product >> 32);
}
/* bits/stl_algo.h:3741-3800; the small-range branch always holds for n <= 65535 */
void mzo_shuffle_perm(mzo_mt19937* g, int n, int* perm) {
    if (n <= 0) return;
    int i = 1;
    if ((n % 2) == 0) {
        int j = (int)mzo_uniform_below(g, 2);
        int t = perm[i]; perm[i] = perm[j]; perm[j] = t;
        ++i;
    }
    while (i != n) {
        uint32_t swap_range = (uint32_t)i + 1U;
        /* __gen_two_uniform_ints(swap_range, swap_range + 1) (:3717-3725) */
        uint32_t b1 = swap_range + 1U;
        uint32_t x = mzo_uniform_below(g, swap_range * b1);
        int p0 = (int)(x / b1), p1 = (int)(x % b1);
        int t = perm[i]; perm[i] = perm[p0]; perm[p0] = t; ++i;
        t = perm[i]; perm[i] = perm[p1]; perm[p1] = t; ++i;
    }
}

/* ------------------------------------------------------------------ gamma / normal / canonical */
/* generate_canonical<double,53> on a 32-bit URBG (bits/random.tcc:3346-3381): two draws */
static double canonical(mzo_mt19937* g) {
    double sum = 0.0, tmp = 1.0;
    sum += (double)mzo_mt_next(g) * tmp; tmp *= 4294967296.0;
    sum += (double)mzo_mt_next(g) * tmp; tmp *= 4294967296.0;
    double ret = sum / tmp;
    if (ret >= 1.0) ret = nextafter(1.0, 0.0);
    return ret;
}
typedef struct { int saved_available; double saved; } normal_state;
/* normal_distribution<double>(0,1)::operator() (bits/random.tcc:1811-1845): Marsaglia polar with a cached spare */
static double normal01(mzo_mt19937* g, normal_state* s) {
    double ret;
    if (s->saved_available) {
        s->saved_available = 0;
        ret = s->saved;
    } else {
        double x, y, r2;
        do {
            x = 2.0 * canonical(g) - 1.0;
            y = 2.0 * canonical(g) - 1.0;
            r2 = x * x + y * y;
        } while (r2 > 1.0 || r2 == 0.0);
        double mult = sqrt(-2 * log(r2) / r2);
        s->saved = x * mult;
        s->saved_available = 1;
        ret = y * mult;
    }
    ret = ret * 1.0 + 0.0;
    return ret;
}
/* gamma_distribution<double>(alpha, 1)::operator() (bits/random.tcc:2336-2393): Marsaglia-Tsang */
static double gamma_draw(mzo_mt19937* g, normal_state* nd, double alpha) {
    double malpha = alpha < 1.0 ? alpha + 1.0 : alpha;
    double a1 = malpha - 1.0 / 3.0;
    double a2 = 1.0 / sqrt(9.0 * a1);
    double u, v, n;
    do {
        do {
            n = normal01(g, nd);
            v = 1.0 + a2 * n;
        } while (v <= 0.0);
        v = v * v * v;
        u = canonical(g);
    } while (u > 1.0 - 0.0331 * n * n * n * n && (log(u) > (0.5 * n * n + a1 * (1.0 - v + log(v)))));
    if (alpha == malpha) return a1 * v * 1.0;
    do { u = canonical(g); } while (u == 0.0);
    return pow(u, 1.0 / alpha) * a1 * v * 1.0;
}
/* util.cpp:28-43.  std::reduce over a vector<double> sums left to right in libstdc++ (bits/stl_numeric.h /
 * numeric: the random-access overload accumulates 4-wide: see reduce4 below). */
static double std_reduce(const double* p, int n) {
    /* libstdc++ 13 <numeric> std::reduce(first,last,init=0.0,plus) for random-access iterators:
     * while (last - first >= 4) { v1 = first[0]+first[1]; v2 = first[2]+first[3]; v3 = v1+v2; init = init+v3; first += 4; }
     * then the tail: init = init + *first. */
    double init = 0.0;
    int i = 0;
    while (n - i >= 4) {
        double v1 = p[i] + p[i + 1];
        double v2 = p[i + 2] + p[i + 3];
        double v3 = v1 + v2;
        init = init + v3;
        i += 4;
    }
    for (; i < n; ++i) init = init + p[i];
    return init;
}
void mzo_sample_dirichlet(mzo_mt19937* g, double alpha, int n, double* out) {
    normal_state nd = {0, 0.0};
    for (int i = 0; i < n; ++i) out[i] = gamma_draw(g, &nd, alpha);
    double sum = std_reduce(out, n);
    for (int i = 0; i < n; ++i) out[i] /= sum;
}

/* ------------------------------------------------------------------ MinMaxStats (mcts.cpp:205-220) */
void mzo_minmax_init(mzo_minmax* m, double minimum, double maximum) {
    m->init_min = minimum; m->init_max = maximum; m->min = minimum; m->max = maximum;
}
void mzo_minmax_reset(mzo_minmax* m) { m->max = m->init_max; m->min = m->init_min; }
void mzo_minmax_update(mzo_minmax* m, double v) {
    m->max = (m->max < v) ? v : m->max; /* std::max(maximum_, value) */
    m->min = (v < m->min) ? v : m->min; /* std::min(minimum_, value) */
}
double mzo_minmax_normalize(const mzo_minmax* m, double v) {
    return (m->max > m->min) ? (v - m->min) / (m->max - m->min) : v;
}

/* ------------------------------------------------------------------ MCTSNode helpers */
static double node_value(const mzo_node* n) { /* mcts.cpp:151-153 */
    return (n->visit_count == 0) ? 0 : n->value_sum / n->visit_count;
}
static void node_init(mzo_node* n, int action, double prior) { /* mcts.h:29 */
    n->visit_count = 0; n->to_play = -1; n->action = action; n->prior = prior;
    n->value_sum = 0; n->reward = 0; n->child_base = -1; n->n_children = 0; n->node_id = -1;
}

mzo_mcts* mzo_mcts_create(const mzo_config* cfg, int32_t seed) { /* mcts.cpp:17-31 */
    mzo_mcts* m = (mzo_mcts*)calloc(1, sizeof(mzo_mcts));
    m->cfg = *cfg;
    mzo_minmax_init(&m->minmax, cfg->value_lowerbound, cfg->value_upperbound);
    mzo_mt_seed(&m->rng, (uint32_t)seed);
    int cap = 1 + cfg->num_actions + (cfg->num_simulations + 8) * cfg->num_actions + 64;
    m->nodes = (mzo_node*)calloc((size_t)cap, sizeof(mzo_node));
    m->n_nodes = 0;
    return m;
}
void mzo_mcts_destroy(mzo_mcts* m) { if (m) { free(m->nodes); free(m); } }

/* MCTSNode::expand (mcts.cpp:157-167) */
void mzo_expand(mzo_mcts* m, int slot, const int32_t* actions, int n_actions, int to_play, double reward,
                const double* policy, int node_id) {
    mzo_node* n = &m->nodes[slot];
    n->to_play = to_play; n->reward = reward; n->node_id = node_id;
    n->child_base = m->n_nodes; n->n_children = n_actions;
    for (int i = 0; i < n_actions; ++i) node_init(&m->nodes[m->n_nodes++], actions[i], policy[actions[i]]);
}
/* MCTSNode::add_exploration_noise (mcts.cpp:170-176) */
void mzo_add_exploration_noise(mzo_mcts* m, int slot, double alpha, double eps) {
    mzo_node* n = &m->nodes[slot];
    double noise[MZO_MAX_ACTIONS];
    mzo_sample_dirichlet(&m->rng, alpha, n->n_children, noise);
    for (int i = 0; i < n->n_children; ++i) {
        mzo_node* c = &m->nodes[n->child_base + i];
        c->prior = c->prior * (1 - eps) + (noise[i] * eps);
    }
}
/* MCTS::ucb_score (mcts.cpp:127-141) */
double mzo_ucb_score(const mzo_mcts* m, const mzo_node* parent, const mzo_node* child) {
    const mzo_config* c = &m->cfg;
    double pb_c = log((parent->visit_count + c->pb_c_base + 1) / c->pb_c_base) + c->pb_c_init;
    pb_c *= sqrt((double)parent->visit_count) / (child->visit_count + 1);
    double prior_score = pb_c * child->prior;
    double value_score = 0;
    if (child->visit_count > 0) {
        double child_value = (c->num_players == 1) ? node_value(child) : -node_value(child);
        value_score = mzo_minmax_normalize(&m->minmax, child->reward + (c->discount * child_value));
    }
    return prior_score + value_score;
}
/* MCTS::select_child (mcts.cpp:110-124): std::shuffle of the child vector, then first strict arg-max */
int mzo_select_child(mzo_mcts* m, int parent_slot) {
    mzo_node* p = &m->nodes[parent_slot];
    int n = p->n_children;
    int perm[MZO_MAX_ACTIONS];
    mzo_node tmp[MZO_MAX_ACTIONS];
    for (int i = 0; i < n; ++i) perm[i] = i;
    mzo_shuffle_perm(&m->rng, n, perm);
    memcpy(tmp, &m->nodes[p->child_base], (size_t)n * sizeof(mzo_node));
    for (int i = 0; i < n; ++i) m->nodes[p->child_base + i] = tmp[perm[i]];
    int selected = -1;
    double best = 0;
    for (int i = 0; i < n; ++i) {
        double v = mzo_ucb_score(m, p, &m->nodes[p->child_base + i]);
        if (selected < 0 || v > best) { selected = p->child_base + i; best = v; }
    }
    return selected;
}
/* MCTS::backpropogate (mcts.cpp:90-107); path_slots[0] is the root, path_slots[len-1] the leaf */
void mzo_backpropagate(mzo_mcts* m, const int32_t* path_slots, int path_len, double value, int to_play) {
    const mzo_config* c = &m->cfg;
    for (int k = path_len - 1; k >= 0; --k) {
        mzo_node* n = &m->nodes[path_slots[k]];
        int same_player = n->to_play == to_play;
        n->value_sum += same_player ? value : -value;
        n->visit_count += 1;
        double new_q = n->reward + (c->discount * (c->num_players == 2 ? -node_value(n) : node_value(n)));
        mzo_minmax_update(&m->minmax, new_q);
        double reward = (same_player && c->num_players == 2) ? -n->reward : n->reward;
        value = reward + (c->discount * value);
    }
}

/* ------------------------------------------------------------------ MCTS::run (mcts.cpp:34-87) */
void mzo_mcts_run(mzo_mcts* m, const int32_t* legal, int n_legal, int to_play, int add_noise,
                  mzo_eval_fn eval, void* ctx, mzo_result* out, int32_t* trace_parent, int32_t* trace_action) {
    const mzo_config* c = &m->cfg;
    int32_t all_actions[MZO_MAX_ACTIONS];
    for (int i = 0; i < c->num_actions; ++i) all_actions[i] = i; /* config.action_space = {0..A-1} (config.h:56-57) */
    mzo_eval_out eo;

    m->n_nodes = 0;
    node_init(&m->nodes[m->n_nodes++], -1, 0); /* root(InvalidAction, 0), mcts.cpp:39 */
    eval(ctx, -1, -1, 0, &eo);
    mzo_expand(m, 0, legal, n_legal, to_play, eo.reward, eo.policy, 0);
    if (add_noise) mzo_add_exploration_noise(m, 0, c->dirichlet_alpha, c->dirichlet_epsilon);
    mzo_minmax_reset(&m->minmax);
    int max_tree_depth = 0;

    int32_t* path = (int32_t*)malloc(sizeof(int32_t) * (size_t)(c->num_simulations + 2));
    for (int i = 0; i < c->num_simulations; ++i) {
        int virtual_to_play = to_play;
        int slot = 0, plen = 0, depth = 0;
        path[plen++] = slot;
        while (m->nodes[slot].n_children > 0) { /* is_expanded(), mcts.cpp:146-148 */
            ++depth;
            slot = mzo_select_child(m, slot);
            path[plen++] = slot;
            virtual_to_play = (virtual_to_play + 1) % c->num_players;
        }
        mzo_node* parent = &m->nodes[path[plen - 2]];
        mzo_node* node = &m->nodes[slot];
        if (trace_parent) trace_parent[i] = parent->node_id;
        if (trace_action) trace_action[i] = node->action;
        eval(ctx, parent->node_id, node->action, i + 1, &eo);
        mzo_expand(m, slot, all_actions, c->num_actions, virtual_to_play, eo.reward, eo.policy, i + 1);
        mzo_backpropagate(m, path, plen, eo.value, virtual_to_play);
        if (depth > max_tree_depth) max_tree_depth = depth;
    }
    free(path);

    /* get_search_statistics (mcts.cpp:179-200) */
    const mzo_node* root = &m->nodes[0];
    int sum_visits = 0;
    memset(out, 0, sizeof(*out));
    out->n_children = root->n_children;
    for (int i = 0; i < root->n_children; ++i) {
        const mzo_node* ch = &m->nodes[root->child_base + i];
        out->child_actions[i] = ch->action;
        out->child_visits[i] = ch->visit_count;
        out->children_relative_visit[i] = ch->visit_count;
        out->emperical_policy[ch->action] = ch->visit_count;
        sum_visits += ch->visit_count;
    }
    for (int i = 0; i < root->n_children; ++i) out->children_relative_visit[i] /= sum_visits;
    for (int i = 0; i < c->num_actions; ++i) out->emperical_policy[i] /= sum_visits;
    out->root_value = node_value(root);
    out->max_tree_depth = max_tree_depth;
}

/* ------------------------------------------------------------------ SE driver (mirrors `ref_driver mcts`) */
typedef struct { const mzo_se_params* p; int na; uint32_t key; uint32_t* cid; } se_ctx;
static void se_eval(void* vctx, int parent_node, int action, int new_node, mzo_eval_out* out) {
    se_ctx* s = (se_ctx*)vctx;
    uint32_t cid = (parent_node < 0) ? se_initial_cid(s->key) : se_child_cid(s->cid[parent_node], (uint32_t)action);
    s->cid[new_node] = cid;
    out->value = se_value(cid, s->p->vlo, s->p->vhi);
    out->reward = se_reward(cid, s->p->rlo, s->p->rhi);
    se_policy(cid, s->na, out->policy);
}
void mzo_run_se(const mzo_config* cfg, const mzo_se_params* se, int seed, int tree, int runs, int legal_mode,
                int add_noise, mzo_result* results, int32_t* trace_parent, int32_t* trace_action) {
    mzo_mcts* m = mzo_mcts_create(cfg, seed + tree); /* one MCTS object per tree slot */
    uint32_t* cid = (uint32_t*)malloc(sizeof(uint32_t) * (size_t)(cfg->num_simulations + 1));
    for (int r = 0; r < runs; ++r) {
        uint32_t key = (uint32_t)(tree * 7919 + r * 104729 + 1) & 0xFFFFFFU;
        se_ctx ctx = {se, cfg->num_actions, key, cid};
        int32_t legal[MZO_MAX_ACTIONS];
        int nl = 0;
        uint32_t mask = se_legal_mask(key, cfg->num_actions, legal_mode);
        for (int a = 0; a < cfg->num_actions; ++a) if (mask & (1U << a)) legal[nl++] = a;
        int to_play = (tree + r) % cfg->num_players;
        mzo_mcts_run(m, legal, nl, to_play, add_noise, se_eval, &ctx, &results[r],
                     trace_parent ? trace_parent + (size_t)r * cfg->num_simulations : NULL,
                     trace_action ? trace_action + (size_t)r * cfg->num_simulations : NULL);
    }
    free(cid);
    mzo_mcts_destroy(m);
}

/* ------------------------------------------------------------------ replay driver (mirrors `ref_driver replay`) */
typedef struct { int na; const double *v, *r, *p; } replay_ctx;
static void replay_eval(void* vctx, int parent_node, int action, int new_node, mzo_eval_out* out) {
    replay_ctx* s = (replay_ctx*)vctx;
    (void)parent_node; (void)action;
    out->value = s->v[new_node];
    out->reward = s->r[new_node];
    memcpy(out->policy, s->p + (size_t)new_node * s->na, sizeof(double) * (size_t)s->na);
}
void mzo_run_replay(const mzo_config* cfg, int seed, uint32_t legal_mask, int to_play, int add_noise,
                    const double* value, const double* reward, const double* policy, mzo_result* result,
                    int32_t* trace_parent, int32_t* trace_action) {
    mzo_mcts* m = mzo_mcts_create(cfg, seed);
    replay_ctx ctx = {cfg->num_actions, value, reward, policy};
    int32_t legal[MZO_MAX_ACTIONS];
    int nl = 0;
    for (int a = 0; a < cfg->num_actions; ++a) if (legal_mask & (1U << a)) legal[nl++] = a;
    mzo_mcts_run(m, legal, nl, to_play, add_noise, replay_eval, &ctx, result, trace_parent, trace_action);
    mzo_mcts_destroy(m);
}

/* ------------------------------------------------------------------ action selection (self_play.cpp:21-39)
 * temperature <= 0: first maximum of the relative visits (std::max_element).  Otherwise weights = visits^(1/T) and one draw
 * of std::discrete_distribution<> (libstdc++ bits/random.tcc: param_type::_M_initialize normalises the weights by their
 * std::accumulate sum, takes partial sums and pins the last to 1; operator() = lower_bound of one generate_canonical<double,53>
 * draw in that table; a single weight draws nothing and returns 0). */
int mzo_select_action(mzo_mt19937* g, const double* rel_visits, const int32_t* actions, int n, double temperature) {
    if (temperature <= 0) {
        int best = 0;
        for (int i = 1; i < n; ++i) if (rel_visits[best] < rel_visits[i]) best = i;
        return actions[best];
    }
    double w[MZO_MAX_ACTIONS], cp[MZO_MAX_ACTIONS];
    for (int i = 0; i < n; ++i) w[i] = pow(rel_visits[i], 1.0 / temperature);
    if (n < 2) return actions[0];
    double sum = 0.0;
    for (int i = 0; i < n; ++i) sum += w[i];
    double run = 0.0;
    for (int i = 0; i < n; ++i) { w[i] /= sum; run += w[i]; cp[i] = run; }
    cp[n - 1] = 1.0;
    const double p = canonical(g);
    int lo = 0, hi = n;   /* std::lower_bound: first element not less than p */
    while (lo < hi) { int mid = lo + (hi - lo) / 2; if (cp[mid] < p) lo = mid + 1; else hi = mid; }
    return actions[lo];
}
