"""Per-move cost of feeding the search: whole stacked observations from the host (mzb_search_batch_slots) versus one new frame per
game into the device-resident ring + on-device stacking (mzb_obs_ring_push + mzb_search_batch_ring).  GPU box."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "muzero-cpp_b200"))
import mzb200
from mzb200.weights import random_weights

game = sys.argv[1] if len(sys.argv) > 1 else "ale"
T = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
sims = int(sys.argv[3]) if len(sys.argv) > 3 else 50
cfg = mzb200.preset(game, max_trees=T, precision=mzb200.BF16, num_simulations=sims)
eng = mzb200.Engine(cfg, 0)
eng.load_weights(random_weights(cfg, 0))
eng.ring_create()
c, h, w = cfg.observation_shape
A = cfg.num_actions
rng = np.random.default_rng(0)
slots = np.arange(T, dtype=np.int32)
masks = np.full(T, (1 << A) - 1, np.uint32); tp = np.zeros(T, np.int32); nz = np.ones(T, np.uint8)
eng.ring_reset(slots)
frames = rng.random((T, c * h * w), dtype=np.float32)
acts = rng.integers(0, A, T).astype(np.int32)
for _ in range(cfg.stacked_observations + 1):
    eng.ring_push(slots, frames, acts)
stacked = eng.ring_stacked(slots)
flat = eng.shapes.initial_flat_size
print("%s: T=%d, frame %.1f KB, stacked observation %.1f KB per game" % (game, T, c * h * w * 4 / 1e3, flat * 4 / 1e3))


def best(fn, n=5):
    ts = []
    for _ in range(n):
        t0 = time.time(); fn(); ts.append(time.time() - t0)
    return min(ts) * 1e3


host_ms = best(lambda: eng.run(stacked, masks, tp, nz, slots=slots))
dev_host = eng.last_search_timing()["total_ms"]
ring_ms = best(lambda: (eng.ring_push(slots, frames, acts), eng.run_ring(slots, masks, tp, nz)))
dev_ring = eng.last_search_timing()["total_ms"]
push_ms = best(lambda: eng.ring_push(slots, frames, acts))
print("move via host-stacked observations : %.2f ms wall (search %.2f ms on the device, H2D %.1f MB)" % (host_ms, dev_host, T * flat * 4 / 1e6))
print("move via ring (push + ring search) : %.2f ms wall (search %.2f ms on the device, push %.2f ms, H2D %.1f MB)" % (ring_ms, dev_ring, push_ms, T * c * h * w * 4 / 1e6))
# the gather kernel alone: stacked bytes written + read once
import ctypes as C
t0 = time.time()
for _ in range(20):
    eng.lib.mzb_obs_ring_stacked  # (symbol check only)
stk_ms = best(lambda: eng.ring_stacked(slots))
print("ring_stacked incl. D2H of %.1f MB: %.2f ms" % (T * flat * 4 / 1e6, stk_ms))
eng.close()
