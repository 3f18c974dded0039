"""CPU suite, part 1: pin the oracle (oracle/) against the reference's own known answers and against golden
vectors produced by the unmodified reference (tests/golden/, generator: tests/golden/make_golden.py)."""
import ctypes as C
import json
import os

import numpy as np
import pytest
import torch

import helpers as H
from oracle import oracle_py as O
from oracle import vprnet_oracle as V


# ---------------------------------------------------------------- search oracle vs the compiled reference
@pytest.mark.parametrize("case", H.load_mcts_golden(), ids=lambda c: c["name"])
def test_oracle_matches_reference_golden(case):
    cfg = H.oracle_config_from(case["config"])
    k = 0
    for t in range(case["trees"]):
        mine = O.run_se(cfg, case["seed"], t, case["runs"], case["legal_mode"], bool(case["noise"]))
        for r in range(case["runs"]):
            ref = case["results"][k]
            k += 1
            assert ref["tree"] == t and ref["run"] == r
            assert H.results_equal(ref, mine[r]), (case["name"], t, r)


@pytest.mark.skipif(not O.have_ref(), reason="oracle/_ref not built")
def test_oracle_matches_live_reference_sweep():
    cfg = O.make_config(num_actions=5, num_simulations=40, discount=0.9)
    txt = O.run_ref("mcts", "game=c4d", "actions=5", "sims=40", "discount=0.9", "trees=24", "runs=2", "legal_mode=1", "seed=77")
    ref = O.parse_ref_mcts(txt)
    for t in range(24):
        mine = O.run_se(cfg, 77, t, 2, 1, True)
        for r in range(2):
            assert H.results_equal(ref[t * 2 + r], mine[r])


# MCTS::run on the reference's own TestEvaluator (tests/mcts_test.cpp:16-27)
TEST_POLICY = [0.01, 0.2, 0.3, 0.1, 0.5, 0.25]


def _test_evaluator(parent, action, new_node):
    return (1.0, 1.0 if parent < 0 else float(action), TEST_POLICY)


def _ref_test_config(**kw):  # tests/mcts_test.cpp:229-241
    base = dict(num_players=2, num_simulations=10, num_actions=6, alpha=0.3, eps=0.25, pb_c_base=1.0, pb_c_init=0.0,
                discount=0.5, lower=0.0, upper=1.0)
    base.update(kw)
    return O.make_config(**base)


APPENDIX_B = [
    ("G1", dict(), 0, [0, 1, 4, 5], False, [
        (2.5531250000000001, 5, [1, 5, 4, 0], [1, 3, 5, 1], [0.1, 0.1, 0, 0, 0.5, 0.3]),
        (2.9281250000000001, 5, [4, 1, 0, 5], [6, 1, 0, 3], [0, 0.1, 0, 0, 0.6, 0.3])]),
    ("G2", dict(), 0, [0, 1, 4, 5], True, [
        (2.859375, 5, [1, 0, 4, 5], [0, 1, 5, 4], [0.1, 0, 0, 0, 0.5, 0.4]),
        (2.8046875, 6, [1, 4, 0, 5], [1, 7, 0, 2], [0, 0.1, 0, 0, 0.7, 0.2])]),
]


@pytest.mark.parametrize("name,kw,seed,legal,noise,expected", APPENDIX_B, ids=[a[0] for a in APPENDIX_B])
def test_appendix_b_goldens(name, kw, seed, legal, noise, expected):
    m = O.mcts_create(_ref_test_config(**kw), seed)
    try:
        for run, (rv, depth, actions, visits, policy) in enumerate(expected):
            r = O.mcts_run(m, legal, run % 2, noise, _test_evaluator)
            assert r["root_value"] == rv
            assert r["depth"] == depth
            assert r["actions"] == actions
            assert r["visits"] == visits
            np.testing.assert_allclose(r["policy"], policy, rtol=0, atol=1e-12)
    finally:
        O.mcts_destroy(m)


def test_appendix_b_g3_g4_defaults():
    # G3: MuZeroConfig defaults, 2 players, bounds +-10, 6 actions, seed 7, all legal, noise on
    m = O.mcts_create(O.make_config(num_actions=6), 7)
    r = O.mcts_run(m, list(range(6)), 0, True, _test_evaluator)
    O.mcts_destroy(m)
    assert r["root_value"] == 1.8830028913384134 and r["depth"] == 8
    assert r["actions"] == [3, 5, 2, 4, 1, 0] and r["visits"] == [0, 13, 7, 26, 4, 0]
    # G4: 1 player, bounds left at the "unknown" sentinels
    m = O.mcts_create(O.make_config(num_actions=6, num_players=1, lower=O.INF_D, upper=O.NINF_D), 7)
    r = O.mcts_run(m, list(range(6)), 0, True, _test_evaluator)
    O.mcts_destroy(m)
    assert r["root_value"] == 16.064277835397309 and r["depth"] == 9
    assert r["actions"] == [1, 2, 5, 4, 0, 3] and r["visits"] == [1, 2, 22, 25, 0, 0]


# ---------------------------------------------------------------- the reference's own unit tests, restated on the oracle
def _expand_chain(m, players):
    """tests/mcts_test.cpp:30-57: expand a left spine of depth 4, rewards 0.1"""
    lib = O.lib()
    actions = (C.c_int32 * 4)(0, 1, 4, 5)
    policy = (C.c_double * 6)(0.01, 0.2, 0.3, 0.1, 0.5, 0.4)
    nodes = m.contents.nodes
    nodes[0].visit_count = 0; nodes[0].action = -1; nodes[0].prior = 0; nodes[0].value_sum = 0
    nodes[0].reward = 0; nodes[0].n_children = 0; nodes[0].child_base = -1
    m.contents.n_nodes = 1
    player, slot, path = 0, 0, [0]
    for _ in range(4):
        lib.mzo_expand(m, slot, actions, 4, player, 0.1, policy, 0)
        player = (player + 1) % players
        slot = nodes[slot].child_base
        nodes[slot].reward = 0.1
        path.append(slot)
    lib.mzo_expand(m, slot, actions, 4, player, 0.1, policy, 0)
    return path, player


@pytest.mark.parametrize("players,expected", [(2, [1, -0.4, 0.3, -0.05, 0.125]), (1, [1, 0.6, 0.4, 0.3, 0.25])])
def test_ref_backpropagate_known_answers(players, expected):  # tests/mcts_test.cpp:30-97
    m = O.mcts_create(_ref_test_config(num_players=players), 0)
    path, player = _expand_chain(m, players)
    arr = (C.c_int32 * len(path))(*path)
    O.lib().mzo_backpropagate(m, arr, len(path), 1.0, player)
    nodes = m.contents.nodes
    for i in range(len(path) - 1, -1, -1):
        n = nodes[path[i]]
        val = 0 if n.visit_count == 0 else n.value_sum / n.visit_count
        assert abs(val - expected[len(expected) - i - 1]) < 1e-3
    O.mcts_destroy(m)


def test_ref_select_child_known_answers():  # tests/mcts_test.cpp:100-135
    lib = O.lib()
    m = O.mcts_create(_ref_test_config(), 0)
    nodes = m.contents.nodes
    m.contents.n_nodes = 1
    nodes[0].visit_count = 0; nodes[0].n_children = 0
    actions = (C.c_int32 * 4)(0, 1, 4, 5)
    policy = (C.c_double * 6)(0.01, 0.2, 0.3, 0.1, 0.5, 0.4)
    lib.mzo_expand(m, 0, actions, 4, 0, 10.0, policy, 0)
    for i in range(4):
        c = nodes[nodes[0].child_base + i]
        c.reward = 1; c.visit_count = 1; c.value_sum = 0.5
    assert nodes[lib.mzo_select_child(m, 0)].action == 0     # root unvisited: the seeded shuffle decides
    nodes[0].visit_count += 1
    assert nodes[lib.mzo_select_child(m, 0)].action == 4     # highest prior
    for i in range(4):
        c = nodes[nodes[0].child_base + i]
        if c.action == 1:
            c.value_sum -= 100; c.visit_count += 100
    assert nodes[lib.mzo_select_child(m, 0)].action == 1
    O.mcts_destroy(m)


def test_ref_ucb_known_answer():  # tests/mcts_test.cpp:138-153
    lib = O.lib()
    m = O.mcts_create(_ref_test_config(), 0)
    nodes = m.contents.nodes
    m.contents.n_nodes = 1
    nodes[0].visit_count = 2; nodes[0].n_children = 0
    actions = (C.c_int32 * 4)(0, 1, 4, 5)
    policy = (C.c_double * 6)(0.01, 0.2, 0.3, 0.1, 0.5, 0.4)
    lib.mzo_expand(m, 0, actions, 4, 0, 10.0, policy, 0)
    c = nodes[nodes[0].child_base]
    c.reward = 2; c.value_sum = -8; c.visit_count = 10
    assert abs(lib.mzo_ucb_score(m, C.byref(nodes[0]), C.byref(c)) - 2.401782) < 1e-3
    O.mcts_destroy(m)


def test_ref_minmax_stats():  # tests/mcts_test.cpp:156-188
    lib = O.lib()
    mm = O.MinMax()
    lib.mzo_minmax_init(C.byref(mm), O.INF_D, O.NINF_D)
    assert lib.mzo_minmax_normalize(C.byref(mm), 10) == 10
    lib.mzo_minmax_update(C.byref(mm), 2)
    assert lib.mzo_minmax_normalize(C.byref(mm), 10) == 10
    lib.mzo_minmax_update(C.byref(mm), 12)
    assert lib.mzo_minmax_normalize(C.byref(mm), 10) == 0.8
    assert lib.mzo_minmax_normalize(C.byref(mm), 2) == 0.0
    assert lib.mzo_minmax_normalize(C.byref(mm), 12) == 1.0
    lib.mzo_minmax_reset(C.byref(mm))
    assert lib.mzo_minmax_normalize(C.byref(mm), 10) == 10
    lib.mzo_minmax_init(C.byref(mm), 0, 1)
    for v in (0.5, 0.0, 1.0):
        assert lib.mzo_minmax_normalize(C.byref(mm), v) == v
    lib.mzo_minmax_update(C.byref(mm), 0.5)
    for v in (0.5, 0.0, 1.0):
        assert lib.mzo_minmax_normalize(C.byref(mm), v) == v


def test_ref_node_expand_and_noise():  # tests/mcts_test.cpp:191-224
    lib = O.lib()
    m = O.mcts_create(_ref_test_config(), 0)
    nodes = m.contents.nodes
    m.contents.n_nodes = 1
    nodes[0].n_children = 0
    acts = [0, 1, 4, 5]
    pol = [0.01, 0.2, 0.3, 0.1, 0.5, 0.25]
    lib.mzo_expand(m, 0, (C.c_int32 * 4)(*acts), 4, 0, 10.0, (C.c_double * 6)(*pol), 0)
    assert nodes[0].n_children == 4 and nodes[0].reward == 10
    for i, a in enumerate(acts):
        c = nodes[nodes[0].child_base + i]
        assert c.prior == pol[a] and c.action == a and c.n_children == 0
    lib.mzo_add_exploration_noise(m, 0, 0.3, 0.25)
    for i, a in enumerate(acts):
        assert nodes[nodes[0].child_base + i].prior != pol[a]
    O.mcts_destroy(m)


def test_shuffle_matches_libstdcxx_golden_order():
    # the all-tie pick of tests/mcts_test.cpp:118-119 (seed 0, 4 children -> action 0) is covered above; here: a
    # permutation is a permutation and consumes the documented number of draws (A=7 -> 3, 4 -> 2, 9 -> 4, 18 -> 9)
    lib = O.lib()
    for n, draws in [(7, 3), (4, 2), (9, 4), (18, 9), (1, 0), (2, 1)]:
        g = O.MT()
        lib.mzo_mt_seed(C.byref(g), 123)
        perm = (C.c_int * n)(*range(n))
        lib.mzo_shuffle_perm(C.byref(g), n, perm)
        assert sorted(perm) == list(range(n))
        consumed = g.idx if g.idx <= 624 and draws else 0
        if draws:
            assert consumed >= draws and consumed <= draws + 3  # rejections are rare


# ---------------------------------------------------------------- ValueEncoder (tests/util_test.cpp)
TRUTH_SCALARS = [[0.0, 1.5, 21.25], [0.0, -10.75, 2.0]]
TRUTH_SUPPORT = [[[0, 0, 0, 0, 1, 0, 0, 0, 0], [0, 0, 0, 0, 0.3896, 0.6104, 0, 0, 0], [0, 0, 0, 0, 0, 0, 0, 0.0834, 0.9166]],
                 [[0, 0, 0, 0, 1, 0, 0, 0, 0], [0, 0.5549, 0.4451, 0, 0, 0, 0, 0, 0], [0, 0, 0, 0, 0.2309, 0.7691, 0, 0, 0]]]


def test_ref_value_encoder_known_answers():  # tests/util_test.cpp:14-100
    enc = V.ValueEncoder(-22, 22)
    assert enc.support_size == 9
    sup = enc.encode(torch.tensor(TRUTH_SCALARS))
    assert torch.allclose(sup, torch.tensor(TRUTH_SUPPORT), 1e-4, 1e-4)
    truth = torch.tensor(TRUTH_SCALARS)
    logits = torch.log(torch.tensor(TRUTH_SUPPORT) + 1e-7)
    for i in range(3):
        assert torch.allclose(enc.decode(logits[:, i, :]), truth[:, i:i + 1], 1e-4, 1e-4)
        assert torch.allclose(enc.decode(torch.tensor(TRUTH_SUPPORT)[:, i, :], False), truth[:, i:i + 1], 1e-4, 1e-4)
    torch.manual_seed(0)
    x = torch.rand(16, 5) * 10
    s = torch.log(enc.encode(x) + 1e-7)
    for i in range(5):
        assert torch.allclose(enc.decode(s[:, i, :]), x[:, i:i + 1], 1e-4, 1e-4)


def test_value_encoder_reference_golden(golden_dir):
    with open(os.path.join(golden_dir, "encoder.json")) as f:
        cases = json.load(f)
    for c in cases:
        enc = V.ValueEncoder(c["min"], c["max"], c["contractive"])
        assert enc.support_size == c["support"]
        for row in c["rows"]:
            logits = torch.tensor([[float.fromhex(x) for x in row["logits"]]], dtype=torch.float32)
            got = float(enc.decode(logits).to(torch.float64)[0, 0])
            assert abs(got - float.fromhex(row["decoded"])) <= 1e-6 * max(1.0, abs(got))


# ---------------------------------------------------------------- network oracle vs VPRNetModel golden outputs
@pytest.mark.parametrize("game", ["tiny", "tinyds"])
def test_network_oracle_matches_reference_golden(game, golden_dir):
    d = O.mzbt_load(os.path.join(golden_dir, "net_%s.mzbt" % game))
    W = {k[2:]: v for k, v in d.items() if k.startswith("w.")}
    net = V.VPRNetOracle(V.preset(game), W)
    o0 = net.initial_inference(d["in.obs"])
    o1 = net.recurrent_inference(o0["hidden"], d["in.a1"])
    o2 = net.recurrent_inference(o1["hidden"], d["in.a2"])
    for pre, o in (("init", o0), ("rec1", o1), ("rec2", o2)):
        for k in ("value", "reward", "policy", "hidden"):
            np.testing.assert_allclose(o[k], d[pre + "." + k], rtol=0, atol=2e-6, err_msg="%s %s %s" % (game, pre, k))
    meta = d["meta"]
    assert net.value_encoder.support_size == meta[7] and net.reward_encoder.support_size == meta[8]
    np.testing.assert_allclose(o0["value_logits"], d["dbg.init.value_logits"], atol=2e-6)
    np.testing.assert_allclose(o1["reward_logits"], d["dbg.rec1.reward_logits"], atol=2e-6)


def test_network_oracle_matches_reference_on_the_shipped_trained_checkpoint(golden_dir):
    """TRAINED weights: tests/golden/checkpoint_c4r_trained.pt is the Connect4 network the reference ships
    (examples/connect4/reanalyze-00/checkpoints/checkpoint--1.pt; its BatchNorm running_var holds denormals ~7e-44, SURVEY §7
    hard part 5), net_c4r_trained.mzbt the unmodified reference's outputs on it (tests/golden/make_golden.py)."""
    import mzb200
    d = O.mzbt_load(os.path.join(golden_dir, "net_c4r_trained.mzbt"))
    W = mzb200.read_checkpoint(os.path.join(golden_dir, "checkpoint_c4r_trained.pt"))
    assert min(float(np.abs(v[v != 0]).min()) for k, v in W.items() if k.endswith("running_var") and (v != 0).any()) < 1e-38
    net = V.VPRNetOracle(V.preset("c4r"), W)
    o0 = net.initial_inference(d["in.obs"])
    o1 = net.recurrent_inference(o0["hidden"], d["in.a1"])
    o2 = net.recurrent_inference(o1["hidden"], d["in.a2"])
    for pre, o in (("init", o0), ("rec1", o1), ("rec2", o2)):
        for k in ("value", "reward", "policy", "hidden"):
            np.testing.assert_allclose(o[k], d[pre + "." + k], rtol=0, atol=5e-6, err_msg="trained %s %s" % (pre, k))


def test_random_weights_have_reference_names(golden_dir):
    for game in ("tiny", "tinyds"):
        d = O.mzbt_load(os.path.join(golden_dir, "net_%s.mzbt" % game))
        ref = {k[2:]: v.shape for k, v in d.items() if k.startswith("w.")}
        mine = {k: v.shape for k, v in H.random_weights(V.preset(game), 0).items()}
        assert ref == mine
