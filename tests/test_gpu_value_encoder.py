"""GPU parity, part 5: ValueEncoder::encode / decode (util.cpp:59-87) through the C ABI (mzb_value_encode / mzb_value_decode)
against the reference's own known answers (tests/util_test.cpp:14-100, the vectors below are its) and against the torch fp32
restatement in oracle/vprnet_oracle.py on random and edge-case inputs."""
import numpy as np
import pytest
import torch

from oracle import vprnet_oracle as V

pytestmark = pytest.mark.gpu

# tests/util_test.cpp: encoder over [-22, 22] (support 9)
TRUTH_SCALARS = [[0.0, 1.5, 21.25], [0.0, -10.75, 2.0]]
TRUTH_SUPPORT = [[[0, 0, 0, 0, 1, 0, 0, 0, 0], [0, 0, 0, 0, 0.3896, 0.6104, 0, 0, 0], [0, 0, 0, 0, 0, 0, 0, 0.0834, 0.9166]],
                 [[0, 0, 0, 0, 1, 0, 0, 0, 0], [0, 0.5549, 0.4451, 0, 0, 0, 0, 0, 0], [0, 0, 0, 0, 0.2309, 0.7691, 0, 0, 0]]]


def _engine(**over):
    import mzb200
    cfg = mzb200.preset("tiny", max_trees=8, precision=0, **over)
    return mzb200.Engine(cfg, 0)


def test_reference_known_answers():
    eng = _engine(min_value=-22.0, max_value=22.0)
    assert eng.shapes.value_support_size == 9
    x = np.array(TRUTH_SCALARS, np.float32).reshape(-1)
    want = np.array(TRUTH_SUPPORT, np.float32).reshape(-1, 9)
    sup = eng.value_encode(x)
    np.testing.assert_allclose(sup, want, rtol=1e-4, atol=1e-4)          # the reference test's own tolerance
    np.testing.assert_allclose(eng.value_decode(np.log(want + 1e-7)), x, rtol=1e-4, atol=1e-4)
    np.testing.assert_allclose(eng.value_decode(want, take_softmax=False), x, rtol=1e-4, atol=1e-4)
    eng.close()


@pytest.mark.parametrize("which,lo,hi", [(0, -10.0, 10.0), (1, -3.0, 3.0), (0, -300.0, 300.0)])
def test_encode_decode_match_the_oracle(which, lo, hi):
    over = dict(min_value=lo, max_value=hi) if which == 0 else dict(min_reward=lo, max_reward=hi)
    eng = _engine(**over)
    enc = V.ValueEncoder(lo, hi)
    S = eng.shapes.value_support_size if which == 0 else eng.shapes.reward_support_size
    assert S == enc.support_size
    rng = np.random.default_rng(3)
    # inside the range, beyond both ends (clamped), the exact ends, zero, and values on support points
    x = np.concatenate([rng.uniform(lo, hi, 4000), rng.uniform(hi, 2 * hi, 50), rng.uniform(2 * lo, lo, 50),
                        [lo, hi, 0.0, -0.0, 1e-8, -1e-8]]).astype(np.float32)
    want = enc.encode(torch.from_numpy(x)).numpy()
    got = eng.value_encode(x, which)
    assert got.shape == want.shape
    np.testing.assert_allclose(got, want, rtol=0, atol=1e-6 * max(2, S))  # fp32: the step index (up to S) differs by one rounding of the division at most
    np.testing.assert_allclose(got.sum(axis=1), 1.0, atol=1e-6)            # a distribution over at most two neighbouring atoms
    assert ((got > 0).sum(axis=1) <= 2).all()
    # decode: the inverse contractive map evaluates sqrt(1 + 4e-3 (|x| + 1.001)) - 1 in fp32 — a cancellation that turns one
    # ulp of the sum over the support into ~1e-4 of the result — so two correct fp32 evaluations (torch's reduction order, the
    # kernel's warp reduction) agree to a few 1e-4 absolute, which is also the tolerance of the reference's own test.  The
    # well-conditioned quantity, h(decoded) = expected support index * step + min, must agree tightly.
    logits = rng.normal(0, 2, (3000, S)).astype(np.float32)
    for got_x, want_x in ((eng.value_decode(logits, which), enc.decode(torch.from_numpy(logits)).numpy()[:, 0]),
                          (eng.value_decode(torch.softmax(torch.from_numpy(logits), 1).numpy(), which, take_softmax=False),
                           enc.decode(torch.softmax(torch.from_numpy(logits), 1), False).numpy()[:, 0])):
        np.testing.assert_allclose(got_x, want_x, rtol=1e-4, atol=5e-4)
        h = lambda v: V._contractive_f32(torch.from_numpy(np.asarray(v, np.float32))).numpy()
        np.testing.assert_allclose(h(got_x), h(want_x), rtol=0, atol=2e-4 * max(1.0, enc.max_value / 4))
    # round trip (tests/util_test.cpp:85-100): decode(log(encode(x) + 1e-7)) == clamp(x)
    xin = x[:4000]
    back = eng.value_decode(np.log(eng.value_encode(xin, which) + 1e-7), which)
    np.testing.assert_allclose(back, xin, rtol=1e-3, atol=1e-3 * max(1.0, hi / 10))
    eng.close()


def test_bad_arguments_fail_loudly():
    import mzb200
    eng = _engine()
    with pytest.raises(mzb200.MzbError):
        eng.value_encode(np.zeros(4, np.float32), which=2)
    eng.close()
