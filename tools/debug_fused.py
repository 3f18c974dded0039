"""Debug: fused bf16 chain vs the fp32 oracle on a small batch (GPU box)."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "muzero-cpp_b200")); sys.path.insert(0, os.path.join(ROOT, "tests"))
import mzb200
import helpers as H
from oracle import vprnet_oracle as V
game = sys.argv[1] if len(sys.argv) > 1 else "c4d"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 13
which = sys.argv[3] if len(sys.argv) > 3 else "both"
ncfg = V.preset(game)
W = H.random_weights(ncfg, seed=7)
net = V.VPRNetOracle(ncfg, W)
eng = mzb200.Engine(mzb200.preset(game, max_trees=max(64, B), precision=mzb200.BF16), 0)
eng.load_weights(W)
rng = np.random.default_rng(1)
obs = rng.random((B, ncfg.in_channels * ncfg.obs_shape[1] * ncfg.obs_shape[2]), dtype=np.float32)
o0 = net.initial_inference(obs)
if which in ("both", "init"):
    g0 = eng.initial_inference(obs)
    for k, attr in (("value", "value"), ("reward", "reward"), ("policy", "policy"), ("hidden", "encoded_state")):
        print("init", k, "max err", float(np.abs(getattr(g0, attr) - o0[k]).max()), "ref max", float(np.abs(o0[k]).max()))
if which in ("both", "rec"):
    a1 = rng.integers(0, ncfg.num_actions, B).astype(np.int32)
    g1 = eng.recurrent_inference(o0["hidden"], a1)
    o1 = net.recurrent_inference(o0["hidden"], a1)
    for k, attr in (("value", "value"), ("reward", "reward"), ("policy", "policy"), ("hidden", "encoded_state")):
        print("rec", k, "max err", float(np.abs(getattr(g1, attr) - o1[k]).max()), "ref max", float(np.abs(o1[k]).max()))
eng.close()
print("done")
