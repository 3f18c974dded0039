"""ALE-shaped initial inference: dump hidden state + value (argv[1] = output npz)"""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "muzero-cpp_b200"))
import mzb200
from mzb200.weights import random_weights
B = 3
cfg = mzb200.preset("ale", max_trees=8, precision=mzb200.BF16)
eng = mzb200.Engine(cfg, 0)
eng.load_weights(random_weights(cfg, 0))
rng = np.random.default_rng(0)
obs = rng.random((B, eng.shapes.initial_flat_size), dtype=np.float32)
g = eng.initial_inference(obs)
np.savez(sys.argv[1], hidden=np.asarray(g.encoded_state), value=np.asarray(g.value), policy=np.asarray(g.policy))
eng.close()
