"""Host-side multi-GPU plumbing: games are sharded over ranks with no per-simulation collective (the
reference's replica model, device_manager.h:68-80, minus the dynamic loans).  One process per GPU; torch.distributed
(NCCL on GPUs, gloo in the CPU tests) carries only the rendezvous, the weight hand-over check and the timing reduction."""
import numpy as np


def shard_games(n_games, rank, world_size):
    """Game / tree-slot g lives on rank g mod world_size (round-robin keeps ragged tails balanced)."""
    return np.arange(rank, n_games, world_size, dtype=np.int64)


def rank_seed(base_seed, game_index):
    """MCTS objects are seeded config.seed + actor number (self_play.cpp:232); the actor number is the GLOBAL game index,
    so a game's search stream does not depend on how many GPUs share the work."""
    return int(base_seed) + int(game_index)


def tensors_checksum(tensors):
    """FNV-1a over the named fp32 tensors in name order (same hash as mzb_weights_checksum uses over the packed blob)."""
    x = 1469598103934665603
    for name in sorted(tensors):
        for b in np.ascontiguousarray(tensors[name], dtype=np.float32).tobytes():
            x = ((x ^ b) * 1099511628211) & 0xFFFFFFFFFFFFFFFF
    return x


def broadcast_tensors(tensors, src, dist):
    """Hand the learner's weights (rank `src`) to every rank through the process group (host path; on GPUs the packed
    device blob goes through one ncclBroadcast instead: mzb_broadcast_weights)."""
    import torch
    names = sorted(tensors)
    out = {}
    for n in names:
        t = torch.from_numpy(np.ascontiguousarray(tensors[n], dtype=np.float32).copy())
        dist.broadcast(t, src=src)
        out[n] = t.numpy()
    return out


def whole_job_rate(units_per_rank, seconds_this_rank, dist=None):
    """Whole-job throughput = units all ranks processed / max over ranks of the time (bench.py contract)."""
    import torch
    if dist is None or not dist.is_initialized():
        return units_per_rank / seconds_this_rank
    t = torch.tensor([seconds_this_rank], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    u = torch.tensor([float(units_per_rank)], dtype=torch.float64)
    dist.all_reduce(u, op=dist.ReduceOp.SUM)
    return float(u[0]) / float(t[0])
