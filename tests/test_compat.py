"""The reference's own class names over libmzb200.so (include/muzero_cpp_mzb200.hpp): muzero_cpp::algorithm::MCTS, Evaluator,
VPRNetEvaluator, DeviceManager, VPRNetModel, MuZeroConfig.  tests/cpp/compat_test.cpp uses them the way tests/mcts_test.cpp,
muzero.cpp and self_play.cpp do.  CPU: the header compiles with plain g++ (no CUDA, torch or abseil headers), links against the
library and fails loudly without a GPU.  GPU: the reference's TestEvaluator searches return the reference's values (SURVEY.md
Appendix B) from the device tree kernels; VPRNetModel matches the reference's network outputs; actor threads over a
VPRNetEvaluator reproduce one direct batched search bit for bit."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "cpp", "compat_test.cpp")
BIN = os.path.join(ROOT, "tests", "cpp", "compat_test")
GOLDEN = os.path.join(ROOT, "tests", "golden")


def _build():
    import mzb200
    mzb200.load_library()   # builds nothing, but fails loudly if the library is absent
    lib = os.path.join(ROOT, "muzero-cpp_b200")
    subprocess.check_call(["g++", "-std=c++17", "-O2", "-Wall", "-I", os.path.join(ROOT, "include"), "-o", BIN, SRC,
                           "-L", lib, "-lmzb200", "-Wl,-rpath," + lib, "-lpthread"])


def test_compat_header_compiles_and_fails_loudly_without_gpu():
    import torch
    _build()
    if torch.cuda.is_available():
        pytest.skip("GPU present: covered by the gpu test")
    out = subprocess.run([BIN, GOLDEN], capture_output=True, text=True, timeout=120)
    assert out.returncode != 0 and "mzb_engine_create" in out.stdout + out.stderr


@pytest.mark.gpu
def test_reference_names_on_the_gpu():
    _build()
    out = subprocess.run([BIN, GOLDEN], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0 and "compat_test: ok" in out.stdout, out.stdout + out.stderr


# ---------------------------------------------------------------- the reference's OWN unit test, unmodified
REF_TEST = "/root/reference/tests/mcts_test.cpp"
REF_BIN = os.path.join(ROOT, "tests", "cpp", "ref_mcts_test")


def build_reference_mcts_test():
    """g++ of /root/reference/tests/mcts_test.cpp AS IT IS against include/compat (the reference's include paths forwarded to
    muzero_cpp_mzb200.hpp) and libmzb200.so.  The source only exists in the build container; the binary travels to the GPU box."""
    lib = os.path.join(ROOT, "muzero-cpp_b200")
    subprocess.check_call(["g++", "-std=c++17", "-O2", "-I", os.path.join(ROOT, "include", "compat"), "-idirafter", "/root/reference",
                           "-o", REF_BIN, REF_TEST, "-L", lib, "-lmzb200", "-Wl,-rpath," + lib, "-lpthread"])


@pytest.mark.skipif(not os.path.exists(REF_TEST), reason="the reference tree is only present in the build container")
def test_reference_mcts_test_compiles_unmodified_against_the_compat_headers():
    import torch
    build_reference_mcts_test()
    if torch.cuda.is_available():
        pytest.skip("GPU present: covered by the gpu test")
    out = subprocess.run([REF_BIN], capture_output=True, text=True, timeout=120)
    assert out.returncode != 0 and "mzb_engine_create" in out.stdout + out.stderr   # no CPU path: fails loudly at the first MCTS object


@pytest.mark.gpu
def test_reference_mcts_test_runs_green_on_the_gpu():
    """tests/mcts_test.cpp of the reference (backpropogate for 2 and 1 players, select_child incl. the seeded all-tie pick,
    ucb_score = 2.401782, MinMaxStats, MCTSNode + add_exploration_noise): exit code 0 = every REQUIRE held."""
    if os.path.exists(REF_TEST):
        build_reference_mcts_test()
    if not os.path.exists(REF_BIN):
        pytest.skip("tests/cpp/ref_mcts_test is built where /root/reference exists (build()); not in this checkout")
    out = subprocess.run([REF_BIN], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stdout + out.stderr


# ---------------------------------------------------------------- cohort drivers vs the reference's play_game / reanalyze
SP_SRC = os.path.join(ROOT, "tests", "cpp", "selfplay_c4_test.cpp")
SP_BIN = os.path.join(ROOT, "tests", "cpp", "selfplay_c4_test")


def _build_selfplay():
    lib = os.path.join(ROOT, "muzero-cpp_b200")
    subprocess.check_call(["g++", "-std=c++17", "-O2", "-Wall", "-I", os.path.join(ROOT, "include"), "-o", SP_BIN, SP_SRC,
                           "-L", lib, "-lmzb200", "-Wl,-rpath," + lib, "-lpthread"])


def test_selfplay_driver_test_compiles():
    _build_selfplay()


@pytest.mark.gpu
def test_cohort_drivers_reproduce_the_reference_self_play_layer():
    """SelfPlayCohort / ReanalyzeCohort vs tests/golden/selfplay_c4.txt (the unmodified play_game / reanalyze over real Connect4,
    128 actor games + their reanalysis + 16 evaluator games): identical histories, bit for bit"""
    _build_selfplay()
    out = subprocess.run([SP_BIN, GOLDEN], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0 and "selfplay_c4_test: ok" in out.stdout, out.stdout[-3000:] + out.stderr[-2000:]
