"""The C++ host facade (include/mzb200_facade.hpp): the reference's blocking one-tree `MCTS::run` call shape on top of the
C ABI.  CPU: it compiles and links against libmzb200.so with plain g++ (no CUDA, no torch headers) and fails loudly
without a GPU.  GPU: T actor threads parked in a Cohort reproduce one direct batched search bit for bit."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "cpp", "facade_test.cpp")
BIN = os.path.join(ROOT, "tests", "cpp", "facade_test")


def _build():
    import mzb200
    mzb200.load_library()   # builds nothing, but fails loudly if the library is absent
    cmd = ["g++", "-std=c++17", "-O2", "-I", os.path.join(ROOT, "include"), "-o", BIN, SRC,
           "-L", os.path.join(ROOT, "muzero-cpp_b200"), "-lmzb200", "-Wl,-rpath," + os.path.join(ROOT, "muzero-cpp_b200"), "-lpthread"]
    subprocess.check_call(cmd)


def test_facade_compiles_and_fails_loudly_without_gpu():
    import torch
    _build()
    if torch.cuda.is_available():
        pytest.skip("GPU present: covered by the gpu test")
    out = subprocess.run([BIN], capture_output=True, text=True, timeout=120)
    assert out.returncode != 0 and "mzb_engine_create" in out.stdout + out.stderr


@pytest.mark.gpu
def test_facade_actor_threads_match_batched_search():
    _build()
    out = subprocess.run([BIN], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0 and "facade_test: ok" in out.stdout, out.stdout + out.stderr
