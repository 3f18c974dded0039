"""Measured bf16-mode error of the network path and what it does to the search, per configuration.

    python tools/bf16_error.py [--out profiles/bf16_error_r02.json] [--positions 1024]

For every config (c4d, c4r, ttt, ale with random-init weights; c4r with the TRAINED checkpoint the reference ships,
tests/golden/checkpoint_c4r_trained.pt) it reports
  * network error of the bf16 engine against the fp32 oracle (oracle/vprnet_oracle.py, pinned on the reference) on the SAME
    inputs: policy max-abs / KL(fp32 || bf16), decoded value and reward absolute error, error in the support-expectation
    domain h(x), and hidden-state error quantiles over live channels — for initial inference and two recurrent steps;
  * decision impact: a 50-simulation search (25 for ttt) over >= 1024 positions with the bf16 engine and with the fp32-check
    engine of this library on identical positions, seeds and noise: how often the most-visited root action differs, the mean
    total-variation distance of the root visit distributions and the mean |root value| difference.
Needs a GPU (both engines are the CUDA library); the oracle is only the checker."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "muzero-cpp_b200"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

import helpers as H  # noqa: E402
import mzb200  # noqa: E402
from oracle import vprnet_oracle as V  # noqa: E402


def h_map(v):
    return np.sign(v) * (np.sqrt(np.abs(v) + 1) - 1) + 0.001 * v


def q(x, ps=(50, 90, 99, 100)):
    x = np.asarray(x, np.float64).ravel()
    return {("p%d" % p if p < 100 else "max"): float(np.percentile(x, p)) for p in ps}


def net_errors(g, o, C):
    pol_g, pol_o = np.asarray(g.policy, np.float64), np.asarray(o["policy"], np.float64)
    kl = (pol_o * (np.log(np.maximum(pol_o, 1e-300)) - np.log(np.maximum(pol_g, 1e-300)))).sum(axis=1)
    out = dict(policy_abs=q(np.abs(pol_g - pol_o)), policy_kl=q(kl),
               policy_argmax_flip=float((pol_g.argmax(axis=1) != pol_o.argmax(axis=1)).mean()))
    for k in ("value", "reward"):
        a, b = np.asarray(getattr(g, k), np.float64), np.asarray(o[k], np.float64)
        out[k + "_abs"] = q(np.abs(a - b))
        out[k + "_h_abs"] = q(np.abs(h_map(a) - h_map(b)))
        out[k + "_range"] = [float(b.min()), float(b.max())]
    got, want = np.asarray(g.encoded_state, np.float64), np.asarray(o["hidden"], np.float64)
    B = want.shape[0]
    raw = np.asarray(o["hidden_raw"], np.float64).reshape(B, C, -1)
    span = raw.max(axis=2) - raw.min(axis=2)
    M = np.abs(raw).max(axis=(1, 2))[:, None]
    live = span > 1e-2 * M                       # channels whose min-max normalisation is well conditioned
    err = np.abs(got - want).reshape(B, C, -1)
    out["hidden_abs_live"] = q(err[live])
    out["hidden_abs_all"] = q(err)
    out["hidden_live_fraction"] = float(live.mean())
    # the tower output itself (before normalisation), relative to its scale: err_raw ~ err_norm * span / M
    out["hidden_rel_to_tower_scale"] = q((err * span[:, :, None] / M[:, :, None])[live])
    return out


def positions(game, n, rng, real):
    """(net config, observation rows, legal masks or None, to_play or None).  real: positions of the stored Connect4 games"""
    ncfg = V.preset(game)
    flat = ncfg.in_channels * ncfg.obs_shape[1] * ncfg.obs_shape[2]
    if real:
        from mzb200 import positions as P
        obs, masks, to_play, _ = P.sample_positions(n, seed=int(rng.integers(0, 1 << 30)))
        return ncfg, obs, masks, to_play
    return ncfg, rng.random((n, flat), dtype=np.float32), None, None


def run_config(name, game, weights, n_net, n_search, binary):
    t0 = time.time()
    rng = np.random.default_rng(7)
    ncfg, obs, _, _ = positions(game, n_net, rng, binary)
    A = ncfg.num_actions
    W = weights(ncfg)
    net = V.VPRNetOracle(ncfg, W)
    res = dict(config=name, game=game, weights="trained (shipped checkpoint)" if name.endswith("trained") else "random-init seed 7",
               net_positions=n_net, inputs="real Connect4 positions of 128 stored games" if binary else "uniform[0,1)")
    eng = mzb200.Engine(mzb200.preset(game, max_trees=max(n_net, 64), precision=mzb200.BF16), 0)
    eng.load_weights(W)
    C = eng.shapes.enc_c
    a1 = rng.integers(0, A, n_net).astype(np.int32)
    a2 = rng.integers(0, A, n_net).astype(np.int32)
    o0 = net.initial_inference(obs)
    res["initial"] = net_errors(eng.initial_inference(obs), o0, C)
    o1 = net.recurrent_inference(o0["hidden"], a1)
    res["recurrent1"] = net_errors(eng.recurrent_inference(o0["hidden"], a1), o1, C)
    o2 = net.recurrent_inference(o1["hidden"], a2)
    res["recurrent2"] = net_errors(eng.recurrent_inference(o1["hidden"], a2), o2, C)
    eng.close()
    # decision impact: bf16 search vs fp32-check search, same positions / seeds / noise
    sims = 25 if game == "ttt" else 50
    _, sobs, masks, to_play = positions(game, n_search, np.random.default_rng(11), binary)
    if masks is None:
        masks = np.full(n_search, (1 << A) - 1, np.uint32)
        to_play = (np.arange(n_search) % (1 if game == "ale" else 2)).astype(np.int32)
    noise = np.ones(n_search, np.uint8)
    out = {}
    for prec in (mzb200.FP32_CHECK, mzb200.BF16):
        e = mzb200.Engine(mzb200.preset(game, max_trees=n_search, precision=prec, num_simulations=sims, seed=3), 0)
        e.load_weights(W)
        out[prec] = e.run(sobs, masks, to_play, noise)
        e.close()
    f, b = out[mzb200.FP32_CHECK], out[mzb200.BF16]
    pf, pb = f.emperical_policy, b.emperical_policy
    res["search"] = dict(positions=n_search, simulations=sims, noise=True,
                         most_visited_action_differs=float((pf.argmax(axis=1) != pb.argmax(axis=1)).mean()),
                         visit_distribution_tv=q(0.5 * np.abs(pf - pb).sum(axis=1), (50, 90, 100)),
                         root_value_abs=q(np.abs(f.root_value - b.root_value), (50, 90, 100)),
                         identical_visit_counts=float((f.child_visits == b.child_visits).all(axis=1).mean()))
    res["seconds"] = round(time.time() - t0, 1)
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(ROOT, "profiles", "bf16_error_r02.json"))
    ap.add_argument("--positions", type=int, default=1024)
    ap.add_argument("--configs", default="c4d,c4r,c4r_real,c4r_trained,ttt,ale")
    a = ap.parse_args()
    rand = lambda ncfg: H.random_weights(ncfg, seed=7)
    trained = lambda ncfg: mzb200.read_checkpoint(os.path.join(ROOT, "tests", "golden", "checkpoint_c4r_trained.pt"))
    table = {
        "c4d": ("c4d", rand, 256, a.positions, False),
        "c4r": ("c4r", rand, 256, a.positions, False),
        "c4r_trained": ("c4r", trained, 256, a.positions, True),       # TRAINED weights on REAL positions
        "c4r_real": ("c4r", rand, 256, a.positions, True),             # random-init weights on the same positions
        "ttt": ("ttt", rand, 256, a.positions, False),
        "ale": ("ale", rand, 24, min(a.positions, 256), False),
    }
    results = []
    for name in a.configs.split(","):
        game, w, n_net, n_search, binary = table[name]
        r = run_config(name, game, w, n_net, n_search, binary)
        results.append(r)
        s = r["search"]
        print("%-12s policy max-abs %.2e KL max %.2e | value abs max %.2e | hidden(live) p99 %.2e max %.2e | most-visited differs %.2f%% (TV p50 %.3f)"
              % (name, r["recurrent1"]["policy_abs"]["max"], r["recurrent1"]["policy_kl"]["max"], r["recurrent1"]["value_abs"]["max"],
                 r["recurrent1"]["hidden_abs_live"]["p99"], r["recurrent1"]["hidden_abs_live"]["max"],
                 100 * s["most_visited_action_differs"], s["visit_distribution_tv"]["p50"]), flush=True)
    with open(a.out, "w") as f:
        json.dump(dict(what="bf16 engine vs fp32 oracle (network) and vs the fp32-check engine (search decisions)", results=results), f, indent=1)
    print("written", a.out)


if __name__ == "__main__":
    main()
