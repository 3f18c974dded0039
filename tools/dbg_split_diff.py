"""recurrent + initial inference on 592 states (148 four-game tiles): dump every output (argv[1] = preset, argv[2] = output npz);
run under different MZB_TOWER_SPLIT settings and compare the files bit for bit"""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "muzero-cpp_b200"))
import mzb200
from mzb200.weights import random_weights
B = 592   # 148 tiles of 4 games
cfg = mzb200.preset(sys.argv[1], max_trees=592, precision=mzb200.BF16)
eng = mzb200.Engine(cfg, 0)
eng.load_weights(random_weights(cfg, 0))
rng = np.random.default_rng(0)
obs = rng.random((B, eng.shapes.initial_flat_size), dtype=np.float32)
g = eng.initial_inference(obs)
act = (np.arange(B) % eng.config.num_actions).astype(np.int32)
r = eng.recurrent_inference(np.asarray(g.encoded_state), act)
np.savez(sys.argv[2], h0=np.asarray(g.encoded_state), v0=np.asarray(g.value), p0=np.asarray(g.policy),
         h1=np.asarray(r.encoded_state), v1=np.asarray(r.value), p1=np.asarray(r.policy), r1=np.asarray(r.reward))
eng.close()
