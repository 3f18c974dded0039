"""Replay sampling and training targets: the CPU restatement (oracle/replay_oracle.py) and the CUDA path
(mzb_replay_*, csrc/replay.cu) against the UNMODIFIED reference — tests/golden/replay_c4.txt, the 128 stored Connect4 histories
pushed through PrioritizedReplayBuffer (capacity 1000, so games are evicted), a sampled Batch every 8 games, priority updates and
sample_game draws, all from one std::mt19937(4242).  Bit-exact: indices, actions, every f64 target and weight, every f32 plane."""
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, os.path.join(ROOT, "muzero-cpp_b200"))
sys.path.insert(0, ROOT)

import helpers as H  # noqa: E402,F401

GOLDEN = os.path.join(HERE, "golden")
FX = float.fromhex


def load_histories():
    from mzb200 import positions as P
    games, cur = [], None
    with open(os.path.join(GOLDEN, "selfplay_c4.txt")) as f:
        for line in f:
            t = line.split()
            if not t:
                continue
            if t[0] == "game":
                cur = {}
                games.append(cur)
            elif cur is None:
                continue
            elif t[0] == "play_actions":
                cur["actions"] = [int(x) for x in t[1:]]
            elif t[0] == "play_rewards":
                cur["rewards"] = [FX(x) for x in t[1:]]
            elif t[0] == "play_to_play":
                cur["to_play"] = [int(x) for x in t[1:]]
            elif t[0] == "play_root_values":
                cur["root_values"] = [FX(x) for x in t[1:]]
            elif t[0] == "play_child_visits":
                v = [FX(x) for x in t[1:]]
                cur["child_visits"] = np.array(v).reshape(-1, 7)
    for g in games:
        obs, _, to_play = P.replay(g["actions"][1:])
        assert to_play == g["to_play"]
        g["observations"] = obs
    return games


def load_events():
    ev, head = [], None
    with open(os.path.join(GOLDEN, "replay_c4.txt")) as f:
        lines = [ln.split() for ln in f if ln.strip()]
    t = lines[0]
    assert t[0] == "replay_c4"
    head = {t[i]: t[i + 1] for i in range(1, len(t), 2)}
    i = 1
    while i < len(lines):
        t = lines[i]
        if t[0] == "save":
            ev.append(("save", int(t[1]), int(t[3]), int(t[5])))
            i += 1
        elif t[0] == "sample":
            b = {}
            for j in range(1, 9):
                r = lines[i + j]
                if r[0] in ("indices", "actions"):
                    b[r[0]] = np.array([int(x) for x in r[1:]], np.int32)
                elif r[0] == "stacked_observations":
                    b[r[0]] = (np.array([float(x) for x in r[1:]]) / 7.0).astype(np.float32)   # printed as value * 7: 0, 7 or the action
                else:
                    b[r[0]] = np.array([FX(x) for x in r[1:]])
            ev.append(("sample", int(t[1]), b))
            i += 9
        elif t[0] == "update":
            ev.append(("update", np.array([FX(x) for x in t[1:]])))
            i += 1
        elif t[0] == "sample_game":
            ev.append(("sample_game", int(t[1]), int(t[2])))
            i += 1
        else:
            raise AssertionError("unknown line " + t[0])
    return head, ev


def check_batch(got, want, what):
    for k in ("indices", "actions"):
        assert np.array_equal(np.asarray(got[k]).ravel(), want[k]), "%s: %s differs" % (what, k)
    for k in ("priorities", "target_values", "target_rewards", "target_policies", "gradient_scale"):
        g = np.asarray(got[k], np.float64).ravel()
        assert g.shape == want[k].shape, "%s: %s shape" % (what, k)
        assert np.array_equal(g.view(np.uint64), want[k].view(np.uint64)), "%s: %s differs (max %g)" % (what, k, np.abs(g - want[k]).max())
    g = np.asarray(got["stacked_observations"], np.float32).ravel()
    assert np.array_equal(g.view(np.uint32), want["stacked_observations"].view(np.uint32)), "%s: stacked_observations differ" % what


def drive(buf, rng, games, events, save):
    """the event stream of the golden file against one implementation; returns how many batches were compared"""
    last, n = None, 0
    for e in events:
        if e[0] == "save":
            save(buf, games[e[1]])
            assert buf.size() == e[2], "size after game %d" % e[1]
            assert int(buf.can_sample()) == e[3]
        elif e[0] == "sample":
            last = buf.sample(rng, e[1])
            check_batch(last, e[2], "batch %d" % n)
            n += 1
        elif e[0] == "update":
            buf.update_history_priorities(last["indices"], e[1])
        elif e[0] == "sample_game":
            assert tuple(buf.sample_game(rng)) == (e[1], e[2])
    return n


def replay_kwargs(head):
    return dict(min_sample_size=int(head["min_sample_size"]), td_steps=int(head["td"]), num_unroll_steps=int(head["unroll"]),
                per_alpha=FX(head["alpha"]), per_beta=FX(head["beta"]), per_epsilon=FX(head["epsilon"]),
                per_beta_increment=FX(head["beta_increment"]))


def test_golden_covers_eviction_and_absorbing_steps():
    head, ev = load_events()
    sizes = [e[2] for e in ev if e[0] == "save"]
    assert max(sizes) > int(head["capacity"])          # SumTree::num_entries_ overshoots while a history is partly overwritten
    assert any(e[0] == "sample_game" and e[2] == 0 for e in ev)      # a history the buffer has lost
    batches = [e[2] for e in ev if e[0] == "sample"]
    assert len(batches) == 16
    tp = np.concatenate([b["target_policies"].reshape(-1, 7) for b in batches])
    assert (np.abs(tp - 1.0 / 7).max(axis=1) == 0).any()             # absorbing steps (uniform policy)


def test_oracle_matches_reference_replay_buffer():
    from oracle import replay_oracle as R
    head, ev = load_events()
    games = load_histories()
    buf = R.ReplayOracle(int(head["capacity"]), (3, 6, 7), 7, action_channels=1, stacked_observations=int(head["stack"]),
                         discount=FX(head["discount"]), **replay_kwargs(head))
    rng = R.StdRng(int(head["seed"]))

    def save(b, g):
        b.save_game_history(R.History(g["observations"], g["actions"], g["rewards"], g["to_play"], g["root_values"], g["child_visits"]))
    assert drive(buf, rng, games, ev, save) == 16


def test_mt19937_callback_source_is_std_mt19937():
    import mzb200
    from oracle import replay_oracle as R
    a, b = mzb200.Mt19937(4242), R.StdRng(4242)
    assert [a() for _ in range(1000)] == [b.next() for _ in range(1000)]


@pytest.mark.gpu
def test_gpu_replay_matches_reference_replay_buffer():
    import mzb200
    head, ev = load_events()
    games = load_histories()
    eng = mzb200.Engine(mzb200.preset("c4d", max_trees=64, stacked_observations=int(head["stack"]), discount=FX(head["discount"]),
                                      evaluator=mzb200.EVAL_SYNTHETIC), 0)
    buf = mzb200.ReplayBuffer(eng, int(head["capacity"]), **replay_kwargs(head))
    rng = mzb200.Mt19937(int(head["seed"]))

    def save(b, g):
        b.save_game_history(g["observations"], g["actions"], g["rewards"], g["to_play"], g["root_values"], g["child_visits"])
    assert drive(buf, rng, games, ev, save) == 16
    assert buf.num_histories() == 128
    buf.close()
    eng.close()


def random_history(rng, L, frame, A, players):
    obs = rng.random((L + 1, frame), dtype=np.float32)
    actions = np.concatenate([[0], rng.integers(0, A, L)]).astype(np.int32)
    rewards = np.concatenate([[0.0], rng.normal(size=L)])
    to_play = (np.arange(L + 1) % players).astype(np.int32) if players > 1 else np.zeros(L + 1, np.int32)
    if players > 1 and L > 4:
        to_play[rng.integers(0, L + 1, 2)] ^= 1          # irregular turn order: the sign rule must follow to_play, not parity
    cv = rng.random((L, A))
    cv /= cv.sum(axis=1, keepdims=True)
    return obs, actions, rewards, to_play, rng.normal(size=L) * 3, cv


@pytest.mark.gpu
@pytest.mark.parametrize("name,capacity,batch,games", [("tinyds", 4096, 64, 60), ("tiny", 257, 32, 40), ("ale", 2000, 8, 6)])
def test_gpu_replay_matches_oracle_on_random_histories(name, capacity, batch, games):
    """other shapes than Connect4: stacked frames with action planes (tinyds 2, ale 6 frames), single player, histories shorter than
    the unroll, a capacity that is not a power of two, td_steps longer than most games — the product against the pinned oracle"""
    import mzb200
    from oracle import replay_oracle as R
    cfg = mzb200.preset(name, max_trees=64, evaluator=mzb200.EVAL_SYNTHETIC)
    eng = mzb200.Engine(cfg, 0)
    c, h, w = cfg.observation_shape
    A, players = cfg.num_actions, cfg.num_players
    kw = dict(min_sample_size=1, td_steps=7, num_unroll_steps=5, per_alpha=0.8, per_beta=0.5, per_epsilon=0.02, per_beta_increment=0.1)
    buf = mzb200.ReplayBuffer(eng, capacity, **kw)
    orc = R.ReplayOracle(capacity, (c, h, w), A, action_channels=cfg.action_channels, stacked_observations=cfg.stacked_observations,
                         discount=cfg.discount, **kw)
    rng = np.random.default_rng(5)
    g_rng, o_rng = mzb200.Mt19937(99), R.StdRng(99)
    compared = 0
    for g in range(games):
        L = int(rng.integers(1, 4)) if g % 5 == 0 else int(rng.integers(5, 120 if name != "ale" else 40))
        hist = random_history(rng, L, c * h * w, A, players)
        buf.save_game_history(*hist)
        orc.save_game_history(R.History(*hist))
        assert buf.size() == orc.size()
        assert buf.total_priority() == orc.tree[0]
        if g % 3 == 2:
            got, want = buf.sample(g_rng, batch), orc.sample(o_rng, batch)
            want = {k: (np.asarray(v).ravel() if k != "num_samples" else v) for k, v in want.items()}
            check_batch(got, want, "%s batch after game %d" % (name, g))
            compared += 1
            err = rng.normal(size=batch) * 2
            buf.update_history_priorities(got["indices"], err)
            orc.update_history_priorities(got["indices"], err)
            assert tuple(buf.sample_game(g_rng)) == tuple(orc.sample_game(o_rng))
    assert compared >= 2
    assert g_rng.draws > 0
    buf.close()
    eng.close()


@pytest.mark.gpu
def test_gpu_replay_update_game_history_and_errors():
    import mzb200
    from oracle import replay_oracle as R
    cfg = mzb200.preset("tiny", max_trees=64, evaluator=mzb200.EVAL_SYNTHETIC)
    eng = mzb200.Engine(cfg, 0)
    c, h, w = cfg.observation_shape
    kw = dict(min_sample_size=1, td_steps=3, num_unroll_steps=2)
    buf = mzb200.ReplayBuffer(eng, 64, **kw)
    orc = R.ReplayOracle(64, (c, h, w), cfg.num_actions, action_channels=cfg.action_channels, stacked_observations=cfg.stacked_observations,
                         discount=cfg.discount, **kw)
    rng = np.random.default_rng(1)
    with pytest.raises(mzb200.MzbError):
        buf.sample(mzb200.Mt19937(0), 4)                 # empty buffer
    hist = random_history(rng, 20, c * h * w, cfg.num_actions, cfg.num_players)
    buf.save_game_history(*hist)
    orc.save_game_history(R.History(*hist))
    # reanalyze write-back (self_play.cpp:216-221): new root values / policies reach the next batch's targets
    rv, cv = rng.normal(size=20), rng.random((20, cfg.num_actions))
    buf.update_game_history(1, rv, cv)
    orc.update_game_history(1, rv, cv)
    buf.update_game_history(7, rv, cv)                   # unknown id: ignored, like SumTree::update_hist
    got, want = buf.sample(mzb200.Mt19937(3), 16), orc.sample(R.StdRng(3), 16)
    check_batch(got, {k: (np.asarray(v).ravel() if k != "num_samples" else v) for k, v in want.items()}, "after update_game_history")
    with pytest.raises(mzb200.MzbError):
        buf.update_history_priorities([5], [1.0])        # not a leaf
    with pytest.raises(mzb200.MzbError):
        buf.save_game_history(*random_history(rng, 100, c * h * w, cfg.num_actions, cfg.num_players))   # longer than the capacity
    buf.close()
    eng.close()
