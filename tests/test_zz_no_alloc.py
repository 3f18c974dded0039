"""mzb_value_decode_static — the allocation-free variant of mzb_value_decode (static __device__ scratch; no counterpart in the
reference) — against the existing entry point, through the C ABI.  Collected last on
purpose.  Same kernel, same stream, same inputs: the outputs must be bit-identical; the comparison
still states the tolerance of the decode tests (1e-4 relative / absolute, tests/test_gpu_value_encoder.py)."""
import ctypes as C
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
for p in (ROOT, os.path.join(ROOT, "muzero-cpp_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)


@pytest.mark.gpu
def test_static_scratch_decode_matches_value_decode_on_two_calls_in_a_row():
    import mzb200
    eng = mzb200.Engine(mzb200.preset("c4d", max_trees=8, precision=mzb200.FP32_CHECK), 0)
    lib, h = eng.lib, eng._h
    fp = C.POINTER(C.c_float)
    rng = np.random.default_rng(17)
    # second call: fewer rows, the other encoder, probabilities instead of logits — the scratch still holds the first call's bytes
    for which, n, softmax in ((0, 3000, 1), (1, 517, 0), (0, 1, 1)):
        S = eng.shapes.value_support_size if which == 0 else eng.shapes.reward_support_size
        lg = rng.normal(0, 2, (n, S)).astype(np.float32)
        if not softmax:
            lg = np.exp(lg) / np.exp(lg).sum(axis=1, keepdims=True)
            lg = lg.astype(np.float32)
        want, got = np.full(n, np.nan, np.float32), np.full(n, np.nan, np.float32)
        assert lib.mzb_value_decode(h, which, lg.ctypes.data_as(fp), n, softmax, want.ctypes.data_as(fp)) == 0, lib.mzb_last_error().decode()
        assert lib.mzb_value_decode_static(h, which, lg.ctypes.data_as(fp), n, softmax, got.ctypes.data_as(fp)) == 0, lib.mzb_last_error().decode()
        assert np.isfinite(got).all()
        np.testing.assert_allclose(got, want, rtol=1e-4, atol=1e-4)
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32))     # same kernel on the same inputs
    # the stated maximum: n * (S + 1) floats; one row more is refused with a message naming the limit, and nothing is written
    S = eng.shapes.value_support_size
    n_max = mzb200.VALUE_DECODE_STATIC_MAX_FLOATS // (S + 1)
    lg = rng.normal(0, 2, (n_max + 1, S)).astype(np.float32)
    out = np.full(n_max + 1, 7.0, np.float32)
    assert lib.mzb_value_decode_static(h, 0, lg.ctypes.data_as(fp), n_max + 1, 1, out.ctypes.data_as(fp)) != 0
    msg = lib.mzb_last_error().decode()
    assert "static" in msg and str(mzb200.VALUE_DECODE_STATIC_MAX_FLOATS) in msg, msg
    assert (out == 7.0).all()
    # exactly the maximum works, and agrees
    got = eng.value_decode_static(lg[:n_max])
    want = eng.value_decode(lg[:n_max])
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    eng.close()
