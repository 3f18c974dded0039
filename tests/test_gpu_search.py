"""GPU parity, part 1: the tree kernels.  Given identical network outputs (the SE synthetic evaluator, evaluated on the
device by the engine and on the CPU by the oracle) and identical seeds, selected actions, visit counts, child order,
root values and tree indices (expansion traces) must match the reference mcts.cpp BIT-EXACTLY.
Called through the C ABI (mzb200.Engine -> libmzb200.so)."""
import numpy as np
import pytest

import helpers as H
from oracle import oracle_py as O

pytestmark = pytest.mark.gpu


def _engine_for(case_cfg, T, seed, **over):
    import mzb200
    kw = dict(case_cfg)
    cfg = mzb200.MuZeroConfig(
        observation_shape=(1, 2, 2), num_actions=kw.get("num_actions", 7), num_players=kw.get("num_players", 2),
        num_simulations=kw.get("num_simulations", 50), discount=kw.get("discount", 0.997),
        dirichlet_alpha=kw.get("alpha", 0.3), dirichlet_epsilon=kw.get("eps", 0.25),
        pb_c_base=kw.get("pb_c_base", 19652.0), pb_c_init=kw.get("pb_c_init", 1.25),
        value_lowerbound={"INF": mzb200.INF_D}.get(kw.get("lower", -10.0), kw.get("lower", -10.0)),
        value_upperbound={"NINF": mzb200.NINF_D}.get(kw.get("upper", 10.0), kw.get("upper", 10.0)),
        seed=seed, max_trees=T, evaluator=mzb200.EVAL_SYNTHETIC,
        network_config=mzb200.MuZeroNetworkConfig(resnet_channels=4), **over)
    return mzb200.Engine(cfg, 0)


def _engine_result(res, t, trace=None):
    L = int(res.n_children[t])
    d = dict(root_value=float(res.root_value[t]), depth=int(res.max_tree_depth[t]),
             actions=[int(x) for x in res.child_actions[t, :L]], visits=[int(x) for x in res.child_visits[t, :L]],
             relvisits=[float(x) for x in res.children_relative_visit[t, :L]],
             policy=[float(x) for x in res.emperical_policy[t]])
    if trace is not None:
        d["trace"] = list(zip([int(x) for x in trace[0][t]], [int(x) for x in trace[1][t]]))
    return d


@pytest.mark.parametrize("case", H.load_mcts_golden(), ids=lambda c: c["name"])
def test_tree_kernels_match_reference_golden(case):
    """golden vectors produced by the unmodified reference (tests/golden/make_golden.py)"""
    T, runs = case["trees"], case["runs"]
    eng = _engine_for(case["config"], T, case["seed"])
    eng.set_recording(True)
    A = eng.config.num_actions
    players = eng.config.num_players
    for r in range(runs):
        keys = H.se_keys(T, r)
        masks = H.se_legal_masks(keys, A, case["legal_mode"])
        to_play = np.array([(t + r) % players for t in range(T)], np.int32)
        res = eng.run(keys.astype(np.float32), masks, to_play, np.full(T, case["noise"], np.uint8))
        trace = eng.get_trace(T)
        for t in range(T):
            ref = case["results"][t * runs + r]
            got = _engine_result(res, t, trace)
            assert H.results_equal(ref, got), (case["name"], t, r, ref, got)
    eng.close()


@pytest.mark.parametrize("A,players,sims,legal_mode", [(7, 2, 50, 0), (9, 2, 25, 1), (18, 1, 50, 0), (3, 1, 64, 1), (32, 2, 20, 1)])
def test_tree_kernels_match_oracle_sweep(A, players, sims, legal_mode):
    """wider seeded sweep against the pinned C oracle, 3 consecutive moves per tree slot"""
    T, runs, seed = 192, 3, 1234
    okw = dict(num_actions=A, num_players=players, num_simulations=sims)
    if players == 1:
        okw.update(lower="INF", upper="NINF")
    eng = _engine_for(okw, T, seed)
    eng.set_recording(True)
    ocfg = H.oracle_config_from(okw)
    oracle = [O.run_se(ocfg, seed, t, runs, legal_mode, True) for t in range(T)]
    for r in range(runs):
        keys = H.se_keys(T, r)
        masks = H.se_legal_masks(keys, A, legal_mode)
        to_play = np.array([(t + r) % players for t in range(T)], np.int32)
        res = eng.run(keys.astype(np.float32), masks, to_play, np.ones(T, np.uint8))
        trace = eng.get_trace(T)
        bad = [t for t in range(T) if not H.results_equal(oracle[t][r], _engine_result(res, t, trace))]
        assert not bad, "run %d: %d trees differ, first %d" % (r, len(bad), bad[0])
    eng.close()


def test_full_size_properties():
    """BASELINE config 2 size (1024 trees, 50 sims, Connect4 shapes): size-independent properties"""
    T = 1024
    eng = _engine_for(dict(), T, 0)
    eng.set_recording(True)
    keys = H.se_keys(T, 0)
    res = eng.run(keys.astype(np.float32), H.se_legal_masks(keys, 7, 0), np.arange(T, dtype=np.int32) % 2, np.ones(T, np.uint8))
    assert (res.child_visits.sum(axis=1) == 50).all()                     # every simulation passes through one root child
    assert np.allclose(res.emperical_policy.sum(axis=1), 1.0)
    assert (np.sort(res.child_actions, axis=1) == np.arange(7)).all()     # shuffled order is a permutation of the legal set
    parent, action = eng.get_trace(T)
    assert (parent[:, 0] == 0).all() and (parent < np.arange(1, 51)[None, :]).all() and (parent >= 0).all()
    # spot-check 16 trees against the oracle at full size
    ocfg = O.make_config()
    for t in range(0, T, 64):
        assert H.results_equal(O.run_se(ocfg, 0, t, 1, 0, True)[0], _engine_result(res, t, (parent, action)))
    tm = eng.last_search_timing()
    assert tm["simulations"] == T * 50 and tm["kernel_launches"] > 100
    eng.close()


def test_edge_cases_and_errors():
    import mzb200
    eng = _engine_for(dict(num_actions=7, num_simulations=5), 4, 0)
    keys = H.se_keys(4, 0).astype(np.float32)
    with pytest.raises(mzb200.MzbError):   # empty legal set: the reference asserts (mcts.cpp:37)
        eng.run(keys, np.array([0, 1, 1, 1], np.uint32), np.zeros(4, np.int32), np.zeros(4, np.uint8))
    with pytest.raises(mzb200.MzbError):   # action outside the action space
        eng.run(keys, np.array([1 << 9, 1, 1, 1], np.uint32), np.zeros(4, np.int32), np.zeros(4, np.uint8))
    with pytest.raises(mzb200.MzbError):   # more trees than slots
        eng.run(np.zeros(8, np.float32), np.ones(8, np.uint32), np.zeros(8, np.int32), np.zeros(8, np.uint8))
    # single legal action: all visits go there, no shuffle draws
    res = eng.run(keys, np.array([1 << 3] * 4, np.uint32), np.zeros(4, np.int32), np.zeros(4, np.uint8))
    assert (res.n_children == 1).all() and (res.child_actions[:, 0] == 3).all() and (res.child_visits[:, 0] == 5).all()
    # ragged batch: fewer trees than slots
    res = eng.run(keys[:2], np.array([0x7f, 0x7f], np.uint32), np.zeros(2, np.int32), np.ones(2, np.uint8))
    assert res.root_value.shape == (2,)
    eng.close()


def _table_evaluator(tree_seed, A):
    """A deterministic evaluator whose outputs depend on the whole (parent node, action, new node) triple — the information a
    user-defined Evaluator subclass (vprnet_evaluator.h:35-41) can see through the encoded states it returns."""
    def ev(parent, action, new_node):
        rng = np.random.default_rng([tree_seed, parent + 1, action + 1, new_node])
        p = rng.random(A) + 0.05
        return float(rng.uniform(-1, 1)), (0.0 if parent < 0 else float(rng.uniform(-0.5, 0.5))), [float(x) for x in p / p.sum()]
    return ev


@pytest.mark.parametrize("A,players,sims,noise", [(6, 2, 10, False), (7, 2, 50, True), (18, 1, 30, True)])
def test_host_evaluator_path_matches_oracle(A, players, sims, noise):
    """MZB_EVAL_HOST: a caller-supplied evaluator behind the device tree kernels (the path muzero_cpp::algorithm::MCTS takes for a
    user Evaluator subclass).  Every tree gets different, history-dependent outputs; decisions must equal the oracle's bit for bit,
    over two consecutive runs on the same RNG streams, with a ragged slot subset in the second."""
    import mzb200
    T, seed = 12, 5
    kw = dict(num_actions=A, num_players=players, num_simulations=sims)
    if players == 1:
        kw.update(lower="INF", upper="NINF")
    eng = _engine_for(kw, T, seed, se_ranges=(0.0, 0.0, 0.0, 0.0))
    eng.close()
    cfg = eng.config
    cfg.evaluator = mzb200.EVAL_HOST
    eng = mzb200.Engine(cfg, 0)
    eng.set_recording(True)
    evs = [_table_evaluator(100 + t, A) for t in range(T)]

    def initial(slots, obs):
        out = [evs[int(s)](-1, -1, 0) for s in slots]
        return [o[0] for o in out], [o[1] for o in out], [o[2] for o in out]

    def recurrent(slots, parent, action, child):
        out = [evs[int(s)](int(p), int(a), child) for s, p, a in zip(slots, parent, action)]
        return [o[0] for o in out], [o[1] for o in out], [o[2] for o in out]

    eng.set_host_evaluator(initial, recurrent)
    okw = dict(kw)
    if players == 1:
        okw.update(lower=O.INF_D, upper=O.NINF_D)
    ocfg = O.make_config(**okw)
    oracles = [O.mcts_create(ocfg, seed + t) for t in range(T)]
    width = eng.shapes.initial_flat_size
    for run, slots in enumerate([list(range(T)), [9, 2, 4, 11, 0]]):
        n = len(slots)
        masks = np.array([((1 << A) - 1) if (s + run) % 3 else (((1 << A) - 1) & ~0b10) for s in slots], np.uint32)
        to_play = np.array([(s + run) % players for s in slots], np.int32)
        res = eng.run(np.zeros((n, width), np.float32), masks, to_play, np.full(n, noise, np.uint8), slots=np.array(slots, np.int32))
        trace = eng.get_trace(n)
        for i, s in enumerate(slots):
            legal = [a for a in range(A) if (int(masks[i]) >> a) & 1]
            want = O.mcts_run(oracles[s], legal, int(to_play[i]), noise, evs[s])
            got = _engine_result(res, i, trace)
            assert H.results_equal(want, got), (run, s, want, got)
    for m in oracles:
        O.mcts_destroy(m)
    # an evaluator that throws surfaces as an error of the search call, not as a silent result
    def broken(slots, parent, action, child):
        raise RuntimeError("boom")
    eng.set_host_evaluator(initial, broken)
    with pytest.raises(mzb200.MzbError):
        eng.run(np.zeros((1, width), np.float32), np.array([1], np.uint32), np.array([0], np.int32), np.zeros(1, np.uint8))
    eng.close()


@pytest.mark.parametrize("shape,channels", [((3, 6, 7), 128), ((1, 3, 3), 5)])
def test_se_state_mover_copies_parent_states(shape, channels):
    """The stub's hidden-state move (se_move_kernel: bulk copies through shared memory; a per-lane loop for odd sizes): after a search
    every expanded node holds its parent's state, hence its tree's root state, and trees do not leak into each other."""
    import mzb200
    T, sims = 37, 24
    cfg = mzb200.MuZeroConfig(observation_shape=shape, num_actions=7, num_players=2, num_simulations=sims, seed=2, max_trees=T,
                              evaluator=mzb200.EVAL_SYNTHETIC, network_config=mzb200.MuZeroNetworkConfig(resnet_channels=channels))
    eng = mzb200.Engine(cfg, 0)
    keys = H.se_keys(T, 0)
    eng.run(keys.astype(np.float32), H.se_legal_masks(keys, 7, 0), np.arange(T, dtype=np.int32) % 2, np.ones(T, np.uint8))
    roots = [eng.se_state(t, 0) for t in range(T)]
    assert roots[0].shape[0] == channels * shape[1] * shape[2]
    assert len({r.tobytes() for r in roots}) > T // 2          # roots differ between trees (position keys differ)
    for t in (0, 1, 17, T - 1):
        assert 0.0 <= roots[t].min() and roots[t].max() < 1.0 and roots[t].std() > 0
        for node in (1, 2, sims // 2, sims):
            assert np.array_equal(eng.se_state(t, node), roots[t]), (t, node)
    eng.close()
