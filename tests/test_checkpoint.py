"""Native reader for the reference's checkpoint files (muzero-cpp_b200/csrc/checkpoint.cpp; SURVEY §8f rank 3, the read half of
VPRNetModel::LoadCheckpoint, vprnet.cpp:66-74).  tests/golden/checkpoint_tiny.pt is the archive the UNMODIFIED reference wrote with
torch::save for the weights of tests/golden/net_tiny.mzbt (tests/golden/make_golden.py).  CPU: the parsed tensors equal those
weights bit for bit, names included; malformed files fail loudly.  GPU: an engine loaded from the file reproduces the reference's
network outputs."""
import os

import numpy as np
import pytest

from oracle import oracle_py as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CKPT = os.path.join(ROOT, "tests", "golden", "checkpoint_tiny.pt")
SHIPPED = "/root/reference/examples/connect4/reanalyze-00/checkpoints/checkpoint--1.pt"


def test_reader_matches_the_reference_weights(golden_dir):
    import mzb200
    want = {k[2:]: v for k, v in O.mzbt_load(os.path.join(golden_dir, "net_tiny.mzbt")).items() if k.startswith("w.")}
    got = mzb200.read_checkpoint(CKPT)
    assert sorted(got) == sorted(want)
    for k in want:
        assert got[k].dtype == np.float32 and got[k].shape == want[k].shape, k
        assert np.array_equal(got[k], want[k]), k


def test_reader_rejects_what_is_not_a_checkpoint(tmp_path):
    import mzb200
    with pytest.raises(mzb200.MzbError, match="cannot open"):
        mzb200.read_checkpoint(str(tmp_path / "missing.pt"))
    junk = tmp_path / "junk.pt"
    junk.write_bytes(b"not a zip archive at all, just some bytes that are long enough")
    with pytest.raises(mzb200.MzbError, match="end-of-central-directory"):
        mzb200.read_checkpoint(str(junk))
    cut = tmp_path / "cut.pt"
    cut.write_bytes(open(CKPT, "rb").read()[:-100])   # central directory damaged
    with pytest.raises(mzb200.MzbError):
        mzb200.read_checkpoint(str(cut))
    data = bytearray(open(CKPT, "rb").read())
    at = data.find(b"\x80\x02c__torch__")
    assert at > 0
    data[at + 2] = 0xEE   # the first opcode of data.pkl (GLOBAL) becomes one no pickle has
    bad = tmp_path / "badop.pt"
    bad.write_bytes(bytes(data))
    with pytest.raises(mzb200.MzbError):
        mzb200.read_checkpoint(str(bad))


@pytest.mark.skipif(not os.path.exists(SHIPPED), reason="the reference tree is only present in the build container")
def test_reader_parses_the_shipped_connect4_checkpoint():
    """the trained Connect4 network the reference ships (README.md:171,194): same tensors as torch.jit.load gives"""
    import torch
    import mzb200
    got = mzb200.read_checkpoint(SHIPPED)
    sd = {k: v.numpy() for k, v in torch.jit.load(SHIPPED, map_location="cpu").state_dict().items() if v.dtype == torch.float32}
    assert sorted(got) == sorted(sd) and len(got) == 118
    for k in sd:
        assert np.array_equal(got[k], sd[k]), k


@pytest.mark.gpu
def test_engine_loaded_from_the_checkpoint_reproduces_the_reference(golden_dir):
    import mzb200
    d = O.mzbt_load(os.path.join(golden_dir, "net_tiny.mzbt"))
    eng = mzb200.Engine(mzb200.preset("tiny", max_trees=8, precision=mzb200.FP32_CHECK), 0)
    eng.load_checkpoint(CKPT)
    g0 = eng.initial_inference(d["in.obs"])
    g1 = eng.recurrent_inference(d["init.hidden"], d["in.a1"])
    for tag, g in (("init", g0), ("rec1", g1)):
        for k, attr in (("value", "value"), ("reward", "reward"), ("policy", "policy"), ("hidden", "encoded_state")):
            np.testing.assert_allclose(getattr(g, attr), d[tag + "." + k], rtol=0, atol=1e-4, err_msg=tag + " " + k)
    with pytest.raises(mzb200.MzbError):
        eng.load_checkpoint(os.path.join(golden_dir, "net_tiny.mzbt"))   # not a module archive
    eng.close()
