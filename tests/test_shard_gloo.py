"""N > 1 host path on CPU: two gloo ranks shard the games, receive rank 0's weights, search their shards with the C
oracle (standing in for the GPU engine, checker role only) and reduce the timing the way bench.py does."""
import os
import sys

import numpy as np
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "muzero-cpp_b200")); sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from mzb200 import shard
    from oracle import oracle_py as O
    n_games = 13   # ragged on purpose
    mine = shard.shard_games(n_games, rank, world)
    # weights: rank 0 draws, everyone receives, checksums agree
    rng = np.random.default_rng(100 + rank)
    W = {"a.weight": rng.random((4, 3), dtype=np.float32), "b.bias": rng.random(5, dtype=np.float32)}
    W = shard.broadcast_tensors(W, 0, dist)
    sums = [None] * world
    dist.all_gather_object(sums, shard.tensors_checksum(W))
    # every rank searches its own games; a game's result depends on its GLOBAL index only
    ocfg = O.make_config()
    res = {int(g): O.run_se(ocfg, 0, shard.rank_seed(0, g), 1, 0, True)[0]["visits"] for g in mine}
    rate = shard.whole_job_rate(len(mine) * 50, 1.0 + rank, dist)   # rank 1 is "slower": max over ranks must win
    allres = [None] * world
    dist.all_gather_object(allres, res)
    if rank == 0:
        q.put(dict(sums=sums, games=sorted(k for r in allres for k in r), rate=rate, res={k: v for r in allres for k, v in r.items()}))
    dist.destroy_process_group()


def test_two_ranks_shard_games_and_agree_on_weights():
    sys.path.insert(0, os.path.join(ROOT, "muzero-cpp_b200"))
    from oracle import oracle_py as O
    O.build_oracle()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 500
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    out = q.get(timeout=180)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert len(set(out["sums"])) == 1                       # replicas hold identical weights
    assert out["games"] == list(range(13))                  # every game searched exactly once
    assert abs(out["rate"] - 13 * 50 / 2.0) < 1e-9          # units of all ranks / max time
    ocfg = O.make_config()                                  # sharding does not change any game's search
    for g in (0, 5, 12):
        assert out["res"][g] == O.run_se(ocfg, 0, g, 1, 0, True)[0]["visits"]
