"""Tower kernel timing (GPU box): recurrent_inference passes back to back + pipelined searches, for A/B runs under env switches.
usage: python tools/tower_time.py [game] [T] [passes] [searches]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "muzero-cpp_b200"))
import mzb200
from mzb200.weights import random_weights

game = sys.argv[1] if len(sys.argv) > 1 else "c4d"
T = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
passes = int(sys.argv[3]) if len(sys.argv) > 3 else 200
searches = int(sys.argv[4]) if len(sys.argv) > 4 else 10
cfg = mzb200.preset(game, max_trees=T, precision=mzb200.BF16)
eng = mzb200.Engine(cfg, 0)
eng.load_weights(random_weights(cfg, 0))
best = 1e9
for _ in range(3):
    ms, nl = eng.bench_recurrent(T, passes)
    best = min(best, ms)
rng = np.random.default_rng(0)
obs = rng.random((T, eng.shapes.initial_flat_size), dtype=np.float32)
masks = np.full(T, (1 << cfg.num_actions) - 1, np.uint32)
tp = (np.arange(T) % cfg.num_players).astype(np.int32)
nz = np.ones(T, np.uint8)
tot, sim = [], []
for i in range(searches + 3):
    eng.run(obs, masks, tp, nz)
    t = eng.last_search_timing()
    if i >= 3:
        tot.append(t["total_ms"]); sim.append(t["sim_loop_ms"])
n = cfg.num_simulations
print("%s T=%d: recurrent %.4f ms/pass (%d launches) = %.3f M states/s | search %.3f ms (sim loop %.3f) = %.3f M sims/s  [env %s]" % (
    game, T, best, nl, T / best / 1e3, np.median(tot), np.median(sim), T * n / np.median(tot) / 1e3,
    " ".join("%s=%s" % (k, v) for k, v in os.environ.items() if k.startswith("MZB_"))))
eng.close()
