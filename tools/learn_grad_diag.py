"""Per-tensor gradient error of mzb_learn_step against the torch-autograd oracle (pinned on the reference), one step.
    python tools/learn_grad_diag.py [game] [batch]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "muzero-cpp_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import helpers as H  # noqa: E402
import mzb200  # noqa: E402
from oracle import oracle_py as O  # noqa: E402
from oracle import vprnet_oracle as V  # noqa: E402

game = sys.argv[1] if len(sys.argv) > 1 else "c4r"
d = O.mzbt_load(os.path.join(ROOT, "tests", "golden", "learn_step_%s.mzbt" % game))
W = {k[2:]: v for k, v in d.items() if k.startswith("w.")} or H.random_weights(V.preset("c4r"), seed=5)
args = (d["in.obs"], d["in.actions"], d["in.target_values"], d["in.target_rewards"], d["in.target_policies"], d["in.priorities"])
net = V.VPRNetOracle(V.preset(game), W)
net.learn_step(*args, d["in.gradient_scale"])
eng = mzb200.Engine(mzb200.preset(game, max_trees=8, precision=mzb200.FP32_CHECK), 0)
eng.load_weights(W)
eng.learn_step(*args, gradient_scale=d["in.gradient_scale"])
rows = []
for k, g in net.grads.items():
    g = g.numpy()
    mine = eng.learn_gradient(k, g.shape)
    rows.append((float(np.abs(mine - g).max()) / max(float(np.abs(g).max()), 1e-12), float(np.abs(g).max()), k))
rows.sort(reverse=True)
for r in rows[:25]:
    print("%.3e  scale %.3e  %s" % r)
print("median rel err %.3e over %d tensors" % (float(np.median([r[0] for r in rows])), len(rows)))
# conditioning: the same oracle with every weight moved by about one fp32 ulp — how far do ITS gradients move?
rng = np.random.default_rng(0)
W2 = {k: (np.asarray(v, np.float32) * (1 + 1.2e-7 * rng.standard_normal(np.asarray(v).shape)).astype(np.float32)) for k, v in W.items()}
net2 = V.VPRNetOracle(V.preset(game), W2)
net2.learn_step(*args, d["in.gradient_scale"])
rows2 = sorted(((float(np.abs(net2.grads[k].numpy() - g.numpy()).max()) / max(float(np.abs(g.numpy()).max()), 1e-12), k) for k, g in net.grads.items()), reverse=True)
for r in rows2[:8]:
    print("oracle vs oracle(+1ulp weights): %.3e  %s" % r)
print("median %.3e" % float(np.median([r[0] for r in rows2])))
