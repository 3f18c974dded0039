"""Profiling target (GPU box): the tree kernels alone — a search with the network stubbed by the SE evaluator — for an ncu
launch list with DRAM byte counters.  Args: game trees sims searches"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "muzero-cpp_b200"))
import mzb200

game = sys.argv[1] if len(sys.argv) > 1 else "c4d"
T = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
sims = int(sys.argv[3]) if len(sys.argv) > 3 else 50
reps = int(sys.argv[4]) if len(sys.argv) > 4 else 2
cfg = mzb200.preset(game, max_trees=T, evaluator=mzb200.EVAL_SYNTHETIC, num_simulations=sims)
eng = mzb200.Engine(cfg, 0)
A = cfg.num_actions
keys = (np.arange(T) + 1).astype(np.float32)
for _ in range(reps):
    eng.run(keys, np.full(T, (1 << A) - 1, np.uint32), (np.arange(T) % cfg.num_players).astype(np.int32), np.ones(T, np.uint8))
    t = eng.last_search_timing()
    print("T=%d sims=%d: sim loop %.3f ms = %.2f us per simulation step, %.2f M sims/s" % (
        T, sims, t["sim_loop_ms"], t["sim_loop_ms"] * 1000 / sims, T * sims / t["sim_loop_ms"] / 1e3))
eng.close()
