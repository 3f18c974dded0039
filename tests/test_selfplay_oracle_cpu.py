"""The self-play helper oracle (oracle/selfplay_oracle.py, oracle/mcts_oracle.c::mzo_select_action) against fixtures written by
the reference build: GameHistory::get_stacked_observations (types.cpp:16-42) run unmodified, and the self_play.cpp:21-39
selection rule over libstdc++'s own std::discrete_distribution / std::mt19937 (tests/golden/make_golden_selfplay.py)."""
import os

import numpy as np

from oracle import selfplay_oracle as SO

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def parse_stack(path):
    cases, cur = [], None
    for line in open(path):
        tok = line.split()
        if tok[0] == "shape":
            cur = dict(kv.split("=") for kv in tok[1:])
            cur = {k: int(v) for k, v in cur.items()}
            cur["steps_out"] = []
            cases.append(cur)
        elif tok[0] == "step":
            cur["steps_out"].append((int(tok[1]), np.array([float.fromhex(x) for x in tok[3:]], np.float32)))
    return cases


def synthetic_history(c, h, w, steps, actions):
    """the game ref_driver `stack` builds (oracle/ref_build/ref_driver.cpp, mode_stack)"""
    k = np.arange(c * h * w, dtype=np.float32)
    obs = [(np.float32(t + 1) + np.float32(0.001) * k).astype(np.float32) for t in range(steps)]
    acts = [0 if t == 0 else (7 * t + 3) % actions for t in range(steps)]
    return obs, acts


def constant_planes(actions, ach, h, w):
    return [np.full(ach * h * w, np.float32(a / actions), np.float32) for a in range(actions)]


def test_stacked_observations_match_the_reference():
    cases = parse_stack(os.path.join(GOLD, "selfplay_stack.txt"))
    assert len(cases) >= 4
    for cs in cases:
        obs, acts = synthetic_history(cs["c"], cs["h"], cs["w"], cs["steps"], cs["actions"])
        planes = constant_planes(cs["actions"], cs["ach"], cs["h"], cs["w"])
        for step, want in cs["steps_out"]:
            got = SO.stacked_observations(obs, acts, step, cs["stack"], (cs["c"], cs["h"], cs["w"]), cs["ach"], planes)
            assert got.shape == want.shape
            assert np.array_equal(got, want), (cs, step)


def parse_select(path):
    runs, cur = [], None
    for line in open(path):
        tok = line.split()
        if tok[0] == "select":
            kv = dict(x.split("=") for x in tok[1:])
            cur = dict(seed=int(kv["seed"]), temperature=float.fromhex(kv["temperature"]), draws=[])
            runs.append(cur)
        elif tok[0] == "draw":
            arrow = tok.index("->")
            pairs = [p.split(":") for p in tok[2:arrow]]
            cur["draws"].append(([int(a) for a, _ in pairs], [float.fromhex(v) for _, v in pairs], int(tok[arrow + 1])))
    return runs


def test_select_action_matches_libstdcxx_discrete_distribution():
    runs = parse_select(os.path.join(GOLD, "selfplay_select.txt"))
    assert len(runs) >= 5
    n = 0
    for r in runs:
        rng = SO.ActorRng(r["seed"])          # one generator per run: the draws consume it in order, like an actor's game
        for actions, rel, want in r["draws"]:
            assert rng.select_action(rel, actions, r["temperature"]) == want, (r["seed"], r["temperature"], actions, rel)
            n += 1
    assert n >= 250
