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


# ---------------------------------------------------------------- the writer (VPRNetModel::SaveCheckpoint, vprnet.cpp:58-63)
def _same_tensors(a, b):
    assert sorted(a) == sorted(b)
    for k in a:
        assert a[k].shape == b[k].shape and np.array_equal(a[k], b[k]), k


@pytest.mark.parametrize("src", ["checkpoint_tiny.pt", "checkpoint_c4r_trained.pt"])
def test_writer_round_trips_through_the_reader_and_through_torch(src, golden_dir, tmp_path):
    """write what the reader parsed -> (a) the library's reader gives the same tensors back bit for bit, (b) torch.jit.load — the
    code path torch::load(model_, path) takes in the reference (vprnet.cpp:71-74) — opens the archive and yields the same state,
    num_batches_tracked included, denormal running_var included"""
    import torch
    import mzb200
    path = os.path.join(golden_dir, src)
    want = mzb200.read_checkpoint(path)
    longs = mzb200.read_checkpoint_longs(path)
    assert longs and all(k.endswith("num_batches_tracked") for k in longs)
    out = str(tmp_path / "checkpoint-7.pt")
    mzb200.write_checkpoint(out, want, longs)
    _same_tensors(mzb200.read_checkpoint(out), want)
    assert mzb200.read_checkpoint_longs(out) == longs
    sd = torch.jit.load(out, map_location="cpu").state_dict()
    ref = torch.jit.load(path, map_location="cpu").state_dict()
    assert list(sd.keys()) == list(ref.keys())          # same modules, same attribute order as the reference's own file
    for k in ref:
        assert sd[k].dtype == ref[k].dtype and torch.equal(sd[k].cpu(), ref[k].cpu()), k


def test_writer_supplies_the_modules_tensors_do_not_name(tmp_path, golden_dir):
    """from float tensors alone: BatchNorm's num_batches_tracked (0) and the MLPs' parameter-less activation modules appear, so
    that torch::load — which reads every registered child and buffer by name — accepts the file"""
    import torch
    import mzb200
    want = mzb200.read_checkpoint(CKPT)
    out = str(tmp_path / "bare.pt")
    mzb200.write_checkpoint(out, want)
    m = torch.jit.load(out, map_location="cpu")
    sd = m.state_dict()
    nbt = [k for k in sd if k.endswith("num_batches_tracked")]
    assert len(nbt) == len([k for k in want if k.endswith("running_var")]) and all(int(sd[k]) == 0 for k in nbt)
    names = [n for n, _ in m.named_modules()]
    assert any(n.endswith("activation_0") for n in names)
    with pytest.raises(mzb200.MzbError):
        mzb200.write_checkpoint(str(tmp_path / "no" / "such" / "dir.pt"), want)


@pytest.mark.skipif(not os.path.exists("/root/reference/muzero-cpp"), reason="needs oracle/_ref/ref_driver (build container)")
def test_reference_loads_what_the_writer_wrote(golden_dir, tmp_path):
    """the UNMODIFIED reference (torch::load in C++, oracle/_ref/ref_driver net ckpt=...) loads the written archive and reproduces
    the golden outputs it produced from its own file"""
    import mzb200
    if not O.have_ref():
        pytest.skip("oracle/_ref not built")
    src = os.path.join(golden_dir, "checkpoint_c4r_trained.pt")
    out = str(tmp_path / "rewritten.pt")
    mzb200.write_checkpoint(out, mzb200.read_checkpoint(src), mzb200.read_checkpoint_longs(src))
    dump, obs_in = str(tmp_path / "net.mzbt"), str(tmp_path / "obs.mzbt")
    want = O.mzbt_load(os.path.join(golden_dir, "net_c4r_trained.mzbt"))
    O.mzbt_save(obs_in, {"in.obs": want["in.obs"]})
    O.run_ref("net", "game=c4r", "seed=5", "no_weights", "ckpt=" + out, "obs_in=" + obs_in, "out=" + dump,
              "path=" + str(tmp_path / "scratch"))
    got = O.mzbt_load(dump)
    for k in want:
        assert np.array_equal(got[k], want[k]), k


@pytest.mark.gpu
def test_engine_saves_what_it_loaded(golden_dir, tmp_path):
    import mzb200
    eng = mzb200.Engine(mzb200.preset("tiny", max_trees=8, precision=mzb200.FP32_CHECK), 0)
    with pytest.raises(mzb200.MzbError):
        eng.save_checkpoint(str(tmp_path / "empty.pt"))   # nothing loaded yet
    eng.load_checkpoint(CKPT)
    out = str(tmp_path / "checkpoint-3.pt")
    eng.save_checkpoint(out)
    _same_tensors(mzb200.read_checkpoint(out), mzb200.read_checkpoint(CKPT))
    assert mzb200.read_checkpoint_longs(out) == mzb200.read_checkpoint_longs(CKPT)
    eng.close()
