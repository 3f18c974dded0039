"""Real Connect4 positions for measurements (BASELINE.json config 4): the action lists of stored games
(tests/golden/selfplay_c4.txt: 128 games the unmodified reference played, see tests/golden/make_golden_selfplay.py) replayed into
observation rows {player-1 stones, player-2 stones, plane of +-1 for the side to move} with the legal-action mask and to_play of
every position — what reanalyze (self_play.cpp:186-225) reads from a GameHistory."""
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
GOLDEN = os.path.join(ROOT, "tests", "golden", "selfplay_c4.txt")


def stored_games(path=GOLDEN):
    """[(actions played, legal masks per step)] for every stored game; actions exclude the leading action_history[0] = 0"""
    games, acts = [], None
    with open(path) as f:
        for line in f:
            tok = line.split()
            if not tok:
                continue
            if tok[0] == "play_actions":
                acts = [int(x) for x in tok[2:]]
            elif tok[0] == "play_legal" and acts is not None:
                games.append((acts, [int(x) for x in tok[1:]]))
                acts = None
    return games


def replay(actions):
    """observation rows [len(actions) + 1, 126], legal masks and to_play of every position of one game"""
    board = np.zeros((6, 7), np.int8)
    mover = 1
    obs, legal, to_play = [], [], []

    def snap():
        o = np.empty((3, 6, 7), np.float32)
        o[0] = board == 1
        o[1] = board == -1
        o[2] = mover
        obs.append(o.reshape(-1))
        legal.append(sum(1 << c for c in range(7) if board[5, c] == 0))
        to_play.append(0 if mover == 1 else 1)
    snap()
    for a in actions:
        r = int(np.argmax(board[:, a] == 0))
        board[r, a] = mover
        mover = -mover
        snap()
    return np.stack(obs), legal, to_play


def sample_positions(n, seed=0, path=GOLDEN):
    """n searched positions drawn uniformly (with replacement) from the stored histories: (obs [n,126], legal mask [n], to_play [n],
    number of histories).  Only positions a search was stored for (not the terminal one) are drawn."""
    games = stored_games(path)
    rows, masks, tps = [], [], []
    for acts, gold_legal in games:
        o, l, t = replay(acts)
        assert l == gold_legal, "replayed legal sets differ from the stored ones"
        rows.append(o[:-1]); masks += l[:-1]; tps += t[:-1]
    rows = np.concatenate(rows)
    rng = np.random.default_rng(seed)
    pick = rng.integers(0, rows.shape[0], n)
    return (np.ascontiguousarray(rows[pick]), np.asarray(masks, np.uint32)[pick], np.asarray(tps, np.int32)[pick], len(games))
