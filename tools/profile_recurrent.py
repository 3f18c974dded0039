"""Profiling target (GPU box): a few recurrent_inference passes + one short search, for ncu launch lists."""
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
sims = int(sys.argv[3]) if len(sys.argv) > 3 else 4
prec = mzb200.BF16 if (len(sys.argv) <= 4 or sys.argv[4] == "bf16") else mzb200.FP32_CHECK
cfg = mzb200.preset(game, max_trees=T, precision=prec, num_simulations=sims)
eng = mzb200.Engine(cfg, 0)
eng.load_weights(random_weights(cfg, 0))
ms, nl = eng.bench_recurrent(T, 2)
print("recurrent: %.3f ms/pass, %d launches, %.0f states/s" % (ms, nl, T / ms * 1e3))
rng = np.random.default_rng(0)
obs = rng.random((T, eng.shapes.initial_flat_size), dtype=np.float32)
reps = int(os.environ.get("MZB_PROFILE_SEARCHES", "1"))
for _ in range(reps):
    res = eng.run(obs, np.full(T, (1 << cfg.num_actions) - 1, np.uint32), np.zeros(T, np.int32), np.ones(T, np.uint8))
print("search:", eng.last_search_timing())
eng.close()
