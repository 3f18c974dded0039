"""root-phase timing experiment: noise on/off (GPU box)"""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "muzero-cpp_b200"))
import mzb200
from mzb200.weights import random_weights
GAME = sys.argv[1] if len(sys.argv) > 1 else "c4d"
T = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
cfg = mzb200.preset(GAME, max_trees=T, precision=mzb200.BF16)
eng = mzb200.Engine(cfg, 0)
eng.load_weights(random_weights(cfg, 0))
rng = np.random.default_rng(0)
obs = rng.random((T, eng.shapes.initial_flat_size), dtype=np.float32)
masks = np.full(T, (1 << cfg.num_actions) - 1, np.uint32); tp = (np.arange(T) % cfg.num_players).astype(np.int32)
for noise in (1, 0, 1, 0):
    nz = np.full(T, noise, np.uint8)
    for _ in range(3):
        t0 = time.time(); eng.run(obs, masks, tp, nz); wall = (time.time() - t0) * 1e3
    print("noise", noise, eng.last_search_timing(), "wall %.2f ms" % wall)
eng.close()
