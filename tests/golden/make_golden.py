"""Regenerate the golden fixtures from the UNMODIFIED reference (oracle/_ref/ref_driver, built from
/root/reference by oracle/ref_build/Makefile).  Run in the build container only:

    make -C oracle/ref_build && python tests/golden/make_golden.py

Outputs (committed):
  mcts_se.json    MCTS::run results + expansion traces for the SE synthetic evaluator over several configs
  net_tiny.mzbt   weights / inputs / outputs of VPRNetModel::{Initial,Recurrent}Inference, plain head
  checkpoint_tiny.pt  the module archive torch::save wrote for net_tiny's weights (what VPRNetModel::LoadCheckpoint reads)
  net_tinyds.mzbt same for the downsample head
  checkpoint_c4r_trained.pt / net_c4r_trained.mzbt  the shipped trained Connect4 checkpoint + the reference's outputs on it
  encoder.json    ValueEncoder support sizes and decode known answers
  learn_step_{tiny,tinyds,c4r}.mzbt  2-3 VPRNetModel::Learn calls on one batch: losses per call, gradients left by the last backward() and
                                 parameters after the last Adam step (c4r: two digests per tensor instead of the tensors)
  learn_{tiny,tinyds,c4r}.mzbt  one VPRNetModel::Learn call: batch, losses, per-sample errors, BatchNorm statistics afterwards
"""
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle import oracle_py as O  # noqa: E402

# (name, ref_driver args, oracle config kwargs, trees, runs, legal_mode, noise, seed)
MCTS_CASES = [
    ("c4_default", ["game=c4d"], dict(), 8, 2, 0, 1, 0),
    ("c4_legal_subsets", ["game=c4d"], dict(), 8, 2, 1, 1, 11),
    ("c4_no_noise", ["game=c4d"], dict(), 4, 2, 1, 0, 3),
    ("ttt_25", ["game=ttt", "sims=25"], dict(num_actions=9, num_simulations=25), 8, 2, 1, 1, 0),
    ("ale_like_1p_unbounded", ["game=tinyds", "actions=18", "unbounded"],
     dict(num_actions=18, num_players=1, lower="INF", upper="NINF"), 4, 2, 0, 1, 5),
    ("ref_test_config", ["game=c4d", "actions=6", "sims=10", "pb_c_base=1", "pb_c_init=0", "discount=0.5", "lower=0", "upper=1"],
     dict(num_actions=6, num_simulations=10, pb_c_base=1.0, pb_c_init=0.0, discount=0.5, lower=0.0, upper=1.0), 4, 2, 1, 1, 0),
    ("deep_4_actions_200", ["game=c4d", "actions=4", "sims=200"], dict(num_actions=4, num_simulations=200), 2, 1, 0, 1, 2),
    ("even_actions_2", ["game=c4d", "actions=2", "sims=30"], dict(num_actions=2, num_simulations=30), 4, 1, 0, 1, 9),
]


def main():
    assert O.have_ref(), "build oracle/_ref first (make -C oracle/ref_build)"
    cases = []
    for name, rargs, okw, T, runs, legal_mode, noise, seed in MCTS_CASES:
        txt = O.run_ref("mcts", *rargs, "trees=%d" % T, "runs=%d" % runs, "legal_mode=%d" % legal_mode,
                        "noise=%d" % noise, "seed=%d" % seed)
        res = O.parse_ref_mcts(txt)
        for r in res:
            r["root_value"] = float(r["root_value"]).hex()
            r["relvisits"] = [float(x).hex() for x in r["relvisits"]]
            r["policy"] = [float(x).hex() for x in r["policy"]]
        cases.append(dict(name=name, config=okw, trees=T, runs=runs, legal_mode=legal_mode, noise=noise, seed=seed, results=res))
    with open(os.path.join(HERE, "mcts_se.json"), "w") as f:
        json.dump(cases, f)
    O.run_ref("net", "game=tiny", "batch=3", "seed=3", "out=" + os.path.join(HERE, "net_tiny.mzbt"),
              "ckpt_out=" + os.path.join(HERE, "checkpoint_tiny.pt"))
    O.run_ref("net", "game=tinyds", "batch=2", "seed=4", "out=" + os.path.join(HERE, "net_tinyds.mzbt"))
    # TRAINED weights: the Connect4 checkpoint the reference ships (README.md:171,194), kept byte for byte as a data fixture;
    # the expected outputs are the unmodified reference's on 32 real Connect4 positions (weights stay in the .pt)
    shipped = "/root/reference/examples/connect4/reanalyze-00/checkpoints/checkpoint--1.pt"
    trained = os.path.join(HERE, "checkpoint_c4r_trained.pt")
    with open(shipped, "rb") as src, open(trained, "wb") as dst:
        dst.write(src.read())
    # inputs: 32 REAL Connect4 positions of the stored games (selfplay_c4.txt; make_golden_selfplay.py writes it first)
    sys.path.insert(0, os.path.join(ROOT, "muzero-cpp_b200"))
    from mzb200 import positions as P
    obs, _, _, _ = P.sample_positions(32, seed=3)
    O.mzbt_save("/tmp/mzb_c4_positions.mzbt", {"in.obs": obs})
    O.run_ref("net", "game=c4r", "seed=5", "no_weights", "ckpt=" + trained, "obs_in=/tmp/mzb_c4_positions.mzbt",
              "out=" + os.path.join(HERE, "net_c4r_trained.mzbt"))
    # VPRNetModel::Learn (vprnet.cpp:192-311): one call on a seeded batch -> batch, LossInfo, BatchNorm buffers after the call
    O.run_ref("learn", "game=tiny", "batch=6", "seed=3", "out=" + os.path.join(HERE, "learn_tiny.mzbt"))
    O.run_ref("learn", "game=tinyds", "batch=3", "seed=4", "out=" + os.path.join(HERE, "learn_tinyds.mzbt"))
    # c4r: the weights are the product generator's (seed 5), handed to the reference through the product's checkpoint WRITER
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import helpers as Hh
    import mzb200
    from oracle import vprnet_oracle as Vv
    # the whole training step (forward, backward, Adam) twice on the same batch: gradients of the last call and parameters after it
    O.run_ref("learn", "game=tiny", "batch=6", "seed=3", "steps=2", "grads=full", "lr=0.01", "l2=0.001", "out=" + os.path.join(HERE, "learn_step_tiny.mzbt"))
    O.run_ref("learn", "game=tinyds", "batch=3", "seed=4", "steps=2", "grads=full", "out=" + os.path.join(HERE, "learn_step_tinyds.mzbt"))
    mzb200.write_checkpoint("/tmp/mzb_learn_c4r_w.pt", Hh.random_weights(Vv.preset("c4r"), seed=5))
    O.run_ref("learn", "game=c4r", "batch=8", "seed=9", "binary_obs=1", "no_weights", "ckpt=/tmp/mzb_learn_c4r_w.pt",
              "post_filter=representation_network.representation_head", "out=" + os.path.join(HERE, "learn_c4r.mzbt"))
    O.run_ref("learn", "game=c4r", "batch=8", "seed=9", "binary_obs=1", "no_weights", "ckpt=/tmp/mzb_learn_c4r_w.pt", "steps=1", "grads=digest",
              "post_filter=representation_network.representation_head", "out=" + os.path.join(HERE, "learn_step_c4r.mzbt"))
    enc = []
    for lo, hi, extra in [(-10, 10, []), (-22, 22, []), (-1, 1, []), (-21, 21, []), (-300, 300, []), (-5, 7, ["no_contractive"])]:
        txt = O.run_ref("encoder", "min=%g" % lo, "max=%g" % hi, "rows=6", "seed=1", *extra)
        lines = txt.strip().splitlines()
        support = int(lines[0].split()[1])
        rows = []
        for ln in lines[1:]:
            tok = ln.split()
            arrow = tok.index("->")
            rows.append(dict(logits=tok[1:arrow], decoded=tok[arrow + 1]))
        enc.append(dict(min=lo, max=hi, contractive=not extra, support=support, rows=rows))
    with open(os.path.join(HERE, "encoder.json"), "w") as f:
        json.dump(enc, f)
    print("golden fixtures written to", HERE)


if __name__ == "__main__":
    main()
