"""Debug helper (GPU box): bf16 tcgen05 backend vs the fp32 oracle, from a single UMMA conv up to the full net."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "muzero-cpp_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import mzb200
from mzb200.weights import random_weights
from oracle import vprnet_oracle as V


def run(game, B, dyn, pred, rep, T=64, seed=0):
    cfg = mzb200.preset(game, max_trees=T, precision=mzb200.BF16)
    cfg.network_config.dynamics_blocks = dyn
    cfg.network_config.prediction_blocks = pred
    cfg.network_config.representation_blocks = rep
    W = random_weights(cfg, seed)
    ncfg = V.preset(game, dyn_blocks=dyn, pred_blocks=pred, rep_blocks=rep)
    net = V.VPRNetOracle(ncfg, W)
    eng = mzb200.Engine(cfg, 0)
    eng.load_weights(W)
    rng = np.random.default_rng(seed)
    hid = rng.random((B, eng.shapes.recurrent_flat_size), dtype=np.float32)
    act = rng.integers(0, cfg.num_actions, B).astype(np.int32)
    g = eng.recurrent_inference(hid, act)
    o = net.recurrent_inference(hid, act)
    out = {}
    for k, attr in (("value", "value"), ("reward", "reward"), ("policy", "policy"), ("hidden", "encoded_state")):
        err = np.abs(getattr(g, attr) - o[k])
        out[k] = (float(err.max()), float(np.abs(o[k]).max()))
    print("%s B=%d dyn=%d pred=%d:" % (game, B, dyn, pred), " ".join("%s err %.4g (scale %.3g)" % (k, v[0], v[1]) for k, v in out.items()), flush=True)
    herr = np.abs(g.encoded_state - o["hidden"]).reshape(B, cfg.network_config.resnet_channels, -1)
    if herr.max() > 0.05:
        b, c, p = np.unravel_index(herr.argmax(), herr.shape)
        print("   worst hidden at sample %d channel %d pixel %d: got %g want %g" % (b, c, p, g.encoded_state.reshape(herr.shape)[b, c, p], o["hidden"].reshape(herr.shape)[b, c, p]))
        print("   per-sample max err:", np.round(herr.max(axis=(1, 2)), 3)[:16])
        print("   per-pixel max err (sample 0):", np.round(herr[0].max(axis=0), 3))
        print("   per-channel max err (sample 0, first 32):", np.round(herr[0].max(axis=1), 3)[:32])
    ob = rng.random((B, eng.shapes.initial_flat_size), dtype=np.float32)
    g0, o0 = eng.initial_inference(ob), net.initial_inference(ob)
    print("   initial: value err %.4g hidden err %.4g policy err %.4g" % (np.abs(g0.value - o0["value"]).max(), np.abs(g0.encoded_state - o0["hidden"]).max(), np.abs(g0.policy - o0["policy"]).max()), flush=True)
    eng.close()


if __name__ == "__main__":
    run("c4d", 5, 0, 0, 0)
    run("c4d", 5, 1, 0, 0)
    run("c4d", 13, 1, 1, 1)
    run("c4d", 64, 4, 4, 4)
    run("c4r", 9, 3, 3, 3)
    run("ttt", 40, 4, 4, 4)
    cfg = mzb200.preset("c4d", max_trees=1024, precision=mzb200.BF16)
    eng = mzb200.Engine(cfg, 0)
    eng.load_weights(random_weights(cfg, 0))
    for B in (1024,):
        ms, nl = eng.bench_recurrent(B, 10)
        print("bench_recurrent B=%d: %.3f ms/pass, %d launches, %.0f states/s" % (B, ms, nl, B / ms * 1e3))
    eng.close()
