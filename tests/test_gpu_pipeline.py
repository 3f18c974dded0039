"""GPU parity, part 4: the pipelined simulation loop (two half-cohorts on two streams, engine_search.cu) must give
bit-identical searches to the single-stream loop — trees are independent and a tree's network arithmetic does not depend on
which other games share its tile.  The switch is an environment variable read once per process, so each arm runs in its
own interpreter."""
import os
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

SCRIPT = r"""
import sys
import numpy as np
ROOT, out, game, T, sims = sys.argv[1], sys.argv[2], sys.argv[3], int(sys.argv[4]), int(sys.argv[5])
sys.path[:0] = [ROOT, ROOT + "/muzero-cpp_b200", ROOT + "/tests"]
import mzb200
import helpers as H
from mzb200.weights import random_weights
cfg = mzb200.preset(game, max_trees=T, precision=mzb200.BF16, num_simulations=sims, seed=9)
eng = mzb200.Engine(cfg, 0)
eng.load_weights(random_weights(cfg, 3))
eng.set_recording(True)
rng = np.random.default_rng(1)
A = cfg.num_actions
keep = {}
for move in range(4):   # move 0 eager, move 1 captured into a CUDA graph, moves 2-3 replay it (the single-stream run has no graph)
    obs = rng.random((T, eng.shapes.initial_flat_size), dtype=np.float32)
    masks = H.se_legal_masks(H.se_keys(T, move), A, 1)
    to_play = rng.integers(0, cfg.num_players, T).astype(np.int32)
    noise = (rng.random(T) < 0.7).astype(np.uint8)
    res = eng.run(obs, masks, to_play, noise)
    parent, action = eng.get_trace(T)
    v, r, p = eng.get_records(T)
    keep.update({"visits%d" % move: res.child_visits, "actions%d" % move: res.child_actions, "root%d" % move: res.root_value,
                 "policy%d" % move: res.emperical_policy, "depth%d" % move: res.max_tree_depth, "parent%d" % move: parent,
                 "action%d" % move: action, "v%d" % move: v, "r%d" % move: r, "p%d" % move: p})
keep["launches"] = np.array(eng.last_search_timing()["kernel_launches"])
np.savez(out, **keep)
eng.close()
"""


def _run(tmp_path, tag, env_value, game, T, sims, **extra_env):
    out = str(tmp_path / ("%s.npz" % tag))
    env = dict(os.environ, MZB_PIPELINE_MIN_TREES=str(env_value), **extra_env)
    proc = subprocess.run([sys.executable, "-c", SCRIPT, ROOT, out, game, str(T), str(sims)], env=env, capture_output=True, text=True, timeout=600)
    assert proc.returncode == 0, proc.stderr[-2000:]
    return np.load(out)


@pytest.mark.parametrize("game,T,sims", [("c4d", 97, 12), ("c4r", 300, 8), ("ale", 40, 6)])
def test_pipelined_loop_is_bit_identical_to_single_stream(tmp_path, game, T, sims):
    single = _run(tmp_path, "single", 0, game, T, sims)       # 0 disables the pipeline
    piped = _run(tmp_path, "piped", 2, game, T, sims)         # every cohort of >= 2 trees is split
    assert int(piped["launches"]) > int(single["launches"]), "the pipelined loop did not run (launch counts equal)"
    for k in single.files:
        if k == "launches":
            continue
        assert np.array_equal(single[k], piped[k]), "pipelined search differs from the single-stream search in %s" % k
    assert (piped["visits1"].sum(axis=1) == sims).all()


@pytest.mark.parametrize("game,T,sims", [("c4d", 64, 10), ("ale", 40, 6)])
def test_staggered_half_tiles_are_bit_identical_to_the_unsplit_tower(tmp_path, game, T, sims):
    """tower_fused_kernel issues the first / last ring slots of a 3x3 layer per half tile (games {0,1}, then {2,3}) so that a half's
    epilogue runs under the other half's MMAs.  The per-column accumulation order is untouched, so whole searches — recorded network
    outputs included — must not change by a bit, whatever the number of slots per half (pipelined cohorts: every tile holds 4 games)."""
    off = _run(tmp_path, "off", 2, game, T, sims, MZB_TOWER_SPLIT="0,0")
    for spec in ("1,1", "3,2"):
        on = _run(tmp_path, "on" + spec.replace(",", "_"), 2, game, T, sims, MZB_TOWER_SPLIT=spec)
        for k in off.files:
            assert np.array_equal(off[k], on[k]), "split %s differs from the unsplit kernel in %s" % (spec, k)
