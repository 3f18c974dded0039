"""GPU parity, part 6: the callers' side — device-resident observation history stacked like GameHistory::get_stacked_observations
(types.cpp:16-42), searches fed from it, and select_action (self_play.cpp:21-39) — against oracle/selfplay_oracle.py (pinned on
the reference in tests/test_selfplay_oracle_cpu.py) and against the plain host-observation search."""
import os

import numpy as np
import pytest

import helpers as H
from oracle import selfplay_oracle as SO
from test_selfplay_oracle_cpu import GOLD, constant_planes, parse_stack, synthetic_history

pytestmark = pytest.mark.gpu


def _engine(game, T, **over):
    import mzb200
    cfg = mzb200.preset(game, max_trees=T, precision=0, **over)
    eng = mzb200.Engine(cfg, 0)
    return eng, cfg


def test_ring_reproduces_the_reference_fixture():
    """the synthetic games of the reference fixture, pushed frame by frame: after every push the ring's stacked observation must be
    the reference's get_stacked_observations(-1, ...) of that game prefix == its output for step t of the full game"""
    for cs in parse_stack(os.path.join(GOLD, "selfplay_stack.txt")):
        if cs["c"] * cs["h"] * cs["w"] == 0:
            continue
        eng, cfg = _engine("tiny", 4, observation_shape=(cs["c"], cs["h"], cs["w"]), stacked_observations=cs["stack"],
                           action_channels=cs["ach"], num_actions=cs["actions"])
        eng.ring_create()                      # NULL planes = constant a / A, what the fixture's game uses
        obs, acts = synthetic_history(cs["c"], cs["h"], cs["w"], cs["steps"], cs["actions"])
        want_by_step = {s: v for s, v in cs["steps_out"]}
        slot = [2]
        eng.ring_reset(slot)
        for t in range(cs["steps"]):
            eng.ring_push(slot, obs[t][None], [acts[t]])
            got = eng.ring_stacked(slot)[0]
            assert np.array_equal(got, want_by_step[t]), (cs, t)
        eng.close()


@pytest.mark.parametrize("game,over", [("tiny", dict()), ("tinyds", dict()), ("c4r", dict(stacked_observations=3, action_channels=1))])
def test_ring_matches_oracle_on_ragged_cohorts(game, over):
    """many slots at different game ages, custom action planes, resets in the middle, pushes for slot subsets"""
    T = 12
    eng, cfg = _engine(game, T, **over)
    c, h, w = cfg.observation_shape
    S, ach, A = cfg.stacked_observations, cfg.action_channels, cfg.num_actions
    rng = np.random.default_rng(4)
    planes = rng.random((A, ach * h * w), dtype=np.float32)
    eng.ring_create(planes)
    hist = {s: ([], []) for s in range(T)}
    eng.ring_reset(np.arange(T))
    for it in range(3 * (S + 2)):
        slots = rng.permutation(T)[: rng.integers(1, T + 1)]
        if it % 5 == 4:                        # some games end: reset a few slots
            dead = slots[: max(1, len(slots) // 3)]
            eng.ring_reset(dead)
            for s in dead:
                hist[int(s)] = ([], [])
        frames = rng.random((len(slots), c * h * w), dtype=np.float32)
        actions = rng.integers(0, A, len(slots)).astype(np.int32)
        eng.ring_push(slots, frames, actions)
        for i, s in enumerate(slots):
            o, a = hist[int(s)]
            a.append(0 if not o else int(actions[i]))
            o.append(frames[i])
        live = np.array([s for s in range(T) if hist[s][0]], np.int32)
        got = eng.ring_stacked(live)
        for i, s in enumerate(live):
            want = SO.stacked_observations(hist[int(s)][0], hist[int(s)][1], -1, S, (c, h, w), ach, planes)
            assert np.array_equal(got[i], want), (game, it, int(s))
    eng.close()


def test_ring_search_equals_host_observation_search():
    """mzb_search_batch_ring == mzb_search_batch_slots on the observations the oracle stacks (same slots, same seeds)"""
    import mzb200
    T, sims = 10, 12
    W = None
    res = {}
    for arm in ("ring", "host"):
        eng, cfg = _engine("tiny", T, num_simulations=sims, seed=21)
        if W is None:
            W = H.random_weights(__import__("oracle.vprnet_oracle", fromlist=["x"]).preset("tiny"), seed=5)
        eng.load_weights(W)
        c, h, w = cfg.observation_shape
        S, ach, A = cfg.stacked_observations, cfg.action_channels, cfg.num_actions
        planes = constant_planes(A, ach, h, w)
        rng = np.random.default_rng(8)
        slots = np.array([7, 1, 4, 9, 0], np.int32)
        eng.ring_create()
        eng.ring_reset(slots)
        hist = {int(s): ([], []) for s in slots}
        out = []
        for move in range(S + 2):
            frames = rng.random((len(slots), c * h * w), dtype=np.float32)
            actions = rng.integers(0, A, len(slots)).astype(np.int32)
            eng.ring_push(slots, frames, actions)
            for i, s in enumerate(slots):
                o, a = hist[int(s)]
                a.append(0 if not o else int(actions[i]))
                o.append(frames[i])
            masks = np.full(len(slots), (1 << A) - 1, np.uint32)
            to_play = ((np.arange(len(slots)) + move) % cfg.num_players).astype(np.int32)
            noise = np.ones(len(slots), np.uint8)
            if arm == "ring":
                r = eng.run_ring(slots, masks, to_play, noise)
            else:
                stacked = np.stack([SO.stacked_observations(hist[int(s)][0], hist[int(s)][1], -1, S, (c, h, w), ach, planes) for s in slots])
                r = eng.run(stacked, masks, to_play, noise, slots=slots)
            out.append((r.child_visits.copy(), r.child_actions.copy(), r.root_value.copy(), r.emperical_policy.copy()))
            assert (r.child_visits.sum(axis=1) == sims).all()
        res[arm] = out
        eng.close()
    for a, b in zip(res["ring"], res["host"]):
        for x, y in zip(a, b):
            assert np.array_equal(x, y)


@pytest.mark.parametrize("temperature", [0.0, 0.25, 1.0])
def test_select_actions_match_the_oracle(temperature):
    """one generator per slot (std::mt19937(seed + actor), self_play.cpp:233), consumed move after move"""
    import mzb200
    T, A = 16, 7
    eng, cfg = _engine("tiny", T, num_actions=A)
    eng.ring_create()
    slots = np.arange(T, dtype=np.int32)[::-1].copy()
    seeds = (100 + slots).astype(np.int32)
    eng.actor_seed(slots, seeds)
    refs = [SO.ActorRng(int(s)) for s in seeds]
    rng = np.random.default_rng(2)
    for move in range(12):
        L = rng.integers(1, A + 1, T)
        visits = np.zeros((T, A), np.int32)
        acts = np.zeros((T, A), np.int32)
        rel = np.zeros((T, A), np.float64)
        for i in range(T):
            acts[i, :L[i]] = rng.permutation(A)[:L[i]]
            visits[i, :L[i]] = rng.integers(0, 30, L[i])
            if visits[i].sum() == 0:
                visits[i, 0] = 1
            if move % 3 == 0 and L[i] > 1:
                visits[i, 1] = visits[i, 0]          # ties: the first maximum wins
            rel[i] = visits[i] / visits[i].sum()     # mcts.cpp:192-194
        stats = mzb200.MCTSReturn(np.zeros(T), np.zeros(T, np.int32), L.astype(np.int32), acts, visits, rel, np.zeros((T, A)))
        got = eng.select_actions(slots, stats, temperature)
        want = [refs[i].select_action(rel[i, :L[i]], acts[i, :L[i]], temperature) for i in range(T)]
        assert got.tolist() == want, (temperature, move)
    eng.close()
