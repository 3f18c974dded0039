"""The forward half of VPRNetModel::Learn (vprnet.cpp:192-311; SURVEY.md 8f rank 2, first slice): K + 1 unrolled TRAIN-mode network
calls + MuzeroNetworkImpl::loss (models.cpp:268-276).  tests/golden/learn_*.mzbt were written by the UNMODIFIED reference
(oracle/_ref/ref_driver learn: one VPRNetModel::Learn call on a seeded types::Batch; tests/golden/make_golden.py): the batch, the
LossInfo it returned, and the BatchNorm running statistics after the call.  CPU: the torch restatement (oracle/vprnet_oracle.py)
is pinned on them.  GPU: mzb_learn_forward through the C ABI reproduces them — the three loss terms and the total to 1e-4, the
per-sample value errors, and the moved running statistics."""
import os

import numpy as np
import pytest

import helpers as H
from oracle import oracle_py as O
from oracle import vprnet_oracle as V

CASES = ["tiny", "tinyds", "c4r"]


def _weights(game, d):
    W = {k[2:]: v for k, v in d.items() if k.startswith("w.")}
    return W if W else H.random_weights(V.preset("c4r"), seed=5)   # c4r: the generator's weights, given to the reference as a checkpoint


def _check(got, d, tol):
    want = d["out.losses"]
    for i, k in enumerate(("total_loss", "value", "policy", "reward")):
        assert abs(got[k] - want[i]) <= tol * max(1.0, abs(want[i])), "%s: %.9g vs the reference's %.9g" % (k, got[k], want[i])
    np.testing.assert_allclose(got["errors"], d["out.errors"], rtol=0, atol=5e-4 * max(1.0, float(np.abs(d["out.errors"]).max())))


@pytest.mark.parametrize("game", CASES)
def test_oracle_learn_forward_matches_the_reference(game, golden_dir):
    d = O.mzbt_load(os.path.join(golden_dir, "learn_%s.mzbt" % game))
    net = V.VPRNetOracle(V.preset(game), _weights(game, d))
    got = net.learn_forward(d["in.obs"], d["in.actions"], d["in.target_values"], d["in.target_rewards"], d["in.target_policies"],
                            d["in.priorities"], value_loss_weight=0.25)
    _check(got, d, 2e-5)
    post = {k[5:]: v for k, v in d.items() if k.startswith("post.")}
    assert post
    for k, v in post.items():   # moved by the forward passes alone: momentum 0.01, unbiased variance
        np.testing.assert_allclose(net.W[k].numpy(), v, rtol=0, atol=2e-6, err_msg=k)


@pytest.mark.gpu
@pytest.mark.parametrize("game", CASES)
def test_learn_forward_matches_the_reference(game, golden_dir):
    import mzb200
    d = O.mzbt_load(os.path.join(golden_dir, "learn_%s.mzbt" % game))
    W = _weights(game, d)
    eng = mzb200.Engine(mzb200.preset(game, max_trees=8, precision=mzb200.FP32_CHECK), 0)
    with pytest.raises(mzb200.MzbError):   # nothing loaded: fails loudly
        eng.learn_forward(d["in.obs"], d["in.actions"], d["in.target_values"], d["in.target_rewards"], d["in.target_policies"], d["in.priorities"])
    eng.load_weights(W)
    got = eng.learn_forward(d["in.obs"], d["in.actions"], d["in.target_values"], d["in.target_rewards"], d["in.target_policies"],
                            d["in.priorities"], gradient_scale=d["in.gradient_scale"], value_loss_weight=0.25)
    _check(got, d, 1e-4)
    for k, v in d.items():
        if k.startswith("post."):
            np.testing.assert_allclose(eng.learn_tensor(k[5:], v.shape), v, rtol=0, atol=1e-5, err_msg=k)
    # a second call starts from the moved statistics (training state persists), the losses do not depend on them
    again = eng.learn_forward(d["in.obs"], d["in.actions"], d["in.target_values"], d["in.target_rewards"], d["in.target_policies"],
                              d["in.priorities"], value_loss_weight=0.25)
    assert abs(again["total_loss"] - got["total_loss"]) <= 1e-9 * max(1.0, abs(got["total_loss"]))
    # inference is untouched by the training-mode passes
    o = V.VPRNetOracle(V.preset(game), W).initial_inference(d["in.obs"])
    np.testing.assert_allclose(eng.initial_inference(d["in.obs"]).policy, o["policy"], rtol=0, atol=1e-4)
    eng.close()
