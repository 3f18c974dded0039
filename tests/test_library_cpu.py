"""CPU suite, part 2: the C-ABI library loads, exports every symbol include/mzb200.h declares, and fails loudly
(no fallback) when there is no CUDA device.  No compute calls here."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_functions():
    with open(os.path.join(ROOT, "include", "mzb200.h")) as f:
        src = f.read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(mzb_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    import mzb200
    lib = mzb200.load_library()
    names = _declared_functions()
    assert len(names) >= 18
    for n in names:
        assert hasattr(lib, n), "libmzb200.so does not export " + n
    assert sorted(mzb200.EXPORTS) == names
    assert b"sm_100a" in lib.mzb_version()


def test_config_struct_matches_header_layout():
    """the ctypes mirror must agree with the C struct: probe through an engine_create that is rejected on validation"""
    import mzb200
    cfg = mzb200.preset("c4d", max_trees=4)
    cfg.num_actions = 99   # invalid -> MZB_ERR_INVALID before any CUDA call
    with pytest.raises(mzb200.MzbError) as ei:
        mzb200.Engine(cfg, 0)
    assert "num_actions" in str(ei.value)
    cfg = mzb200.preset("c4d", max_trees=4)
    cfg.num_players = 3
    with pytest.raises(mzb200.MzbError) as ei:
        mzb200.Engine(cfg, 0)
    assert "player" in str(ei.value)


def test_no_silent_fallback_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    import mzb200
    with pytest.raises(mzb200.MzbError):
        mzb200.Engine(mzb200.preset("c4d", max_trees=4), 0)


def test_missing_library_fails_loudly(tmp_path, monkeypatch):
    import mzb200
    monkeypatch.setattr(mzb200, "_lib", None)
    with pytest.raises(mzb200.MzbError):
        mzb200.load_library(str(tmp_path / "nope.so"))
    monkeypatch.setattr(mzb200, "_lib", None)
    mzb200.load_library()


def test_product_never_imports_oracle():
    """the product path must not import, include, link or call anything under oracle/ (checker only)"""
    pkg = os.path.join(ROOT, "muzero-cpp_b200")
    bad = re.compile(r"from\s+oracle|import\s+oracle|oracle/|oracle_py|vprnet_oracle|mzo_|libmzb_oracle")
    for dp, _, files in os.walk(pkg):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cpp", ".h", ".cuh")):
                with open(os.path.join(dp, fn)) as f:
                    assert not bad.search(f.read()), fn
