#!/usr/bin/env python
"""bench.py — MCTS simulations/s of the self-play search hot path (BASELINE.json metric) on N B200s.

A "step" is one pass of the hot path over one batch: T concurrent searches (initial inference + num_simulations x
{tree step, recurrent inference} + statistics) on synthetic observations of the workload's shape, random-init weights.
  value : whole-job sims/s with the observations already resident in HBM (device time, CUDA events, max over ranks)
  e2e   : same metric through the public C-ABI call with HOST buffers (H2D of observations + D2H of results inside)
  roofline     : the residual-tower kernels (tensor bound): F_rec x states/s vs the measured bf16 peak
  roofline_tree: the tree kernels alone (HBM bound; network stubbed by the SE evaluator)
  cpu_baseline : the reference's own CPU path (oracle/_ref/ref_driver bench) on this box's host cores, bounded sample
`--impl reference` times only that CPU path and prints the same line shape.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "muzero-cpp_b200"))

METRIC = "mcts_sims_per_s"
UNIT = "sims/s"
WORKLOADS = {  # BASELINE.json configs that fit one GPU; configs[1] is the headline
    "c4d": dict(desc="Connect4 example, 50 sims/move, ResNet tower 128ch x 4+4+4 blocks, 1024 concurrent games per GPU", trees=1024),
    "c4r": dict(desc="Connect4, README/shipped tower 64ch x 3 blocks, 50 sims, 1024 concurrent games per GPU", trees=1024),
    "reanalyze": dict(desc="Connect4 reanalyze: batched re-search of 4096 replay positions per GPU (drawn from 128 stored games, "
                           "legal sets and to_play from the history, through the observation ring), 50 sims", trees=4096, preset="c4d", ring=True),
    "ale": dict(desc="ALE/Pong-shaped synthetic 96x96 stacked frames, downsample ResNet, 50 sims, 18 actions, 1024 games per GPU", trees=1024),
    "ttt": dict(desc="TicTacToe, 25 sims/move (reference CPU-runnable case)", trees=1024),
}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return dict(hbm=d["hbm_gbs"], tf_burst=d["bf16_tflops"], tf_sustained=d["bf16_tflops_sustained"], source="measured")
    return dict(hbm=6650.0, tf_burst=1590.0, tf_sustained=1400.0, source="fallback")


def ncu_traffic(workload, trees):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the dominant kernel, from the committed ncu --set full
    summary (profiles/); None when no capture exists for this workload and cohort size."""
    p = os.path.join(ROOT, "profiles", "tower_fused_traffic.json")
    try:
        with open(p) as f:
            d = json.load(f)
        return d.get("%s_T%d" % (workload, trees), {}).get("dram_bytes_per_launch")
    except Exception:
        return None


class ClockSampler:
    """SM clock and throttle reasons sampled DURING the timed region (B200_PROFILING.md clocks line).  NVML in a
    thread every 25 ms (NVML queries take driver locks that delay kernel launches: sampled sparsely on purpose; the default 20 steps give ~10 samples); falls back to
    the nvidia-smi loop when the NVML binding is missing."""
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
    BITS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap"}

    def __init__(self, index):
        self.index, self.samples, self.proc, self.nvml, self.stop_flag = index, [], None, None, False
        self.sm, self.mx, self.reasons = [], None, set()

    def start(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            self.h = pynvml.nvmlDeviceGetHandleByIndex(self.index)
            self.mx = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self.nvml = pynvml
            self.thread = threading.Thread(target=self._poll, daemon=True)
            self.thread.start()
            return
        except Exception:
            self.nvml = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _poll(self):
        n = self.nvml
        while not self.stop_flag:
            try:
                self.sm.append(float(n.nvmlDeviceGetClockInfo(self.h, n.NVML_CLOCK_SM)))
                r = n.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in self.BITS.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.025)

    def _read(self):
        for line in self.proc.stdout:
            self.samples.append(line.strip())

    def stop(self):
        if self.nvml:
            self.stop_flag = True
            self.thread.join(timeout=1)
            sm = sorted(self.sm)
            return dict(sm_mhz=sm[len(sm) // 2] if sm else None, sm_max_mhz=self.mx, reasons=sorted(self.reasons), samples=len(sm), source="nvml")
        if not self.proc:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=["nvidia-smi unavailable"])
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for s in self.samples:
            tok = [t.strip() for t in s.split(",")]
            if len(tok) < 7:
                continue
            try:
                sm.append(float(tok[0])); mx = float(tok[1])
            except ValueError:
                continue
            for n, v in zip(names, tok[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return dict(sm_mhz=sm[len(sm) // 2] if sm else None, sm_max_mhz=mx, reasons=sorted(reasons), samples=len(sm), source="nvidia-smi")


def run_reference_cpu(game, seconds_target, sims, steps=3, warmup=1):
    """The reference's own CPU path on this box's host cores (oracle/_ref/ref_driver bench: unmodified MCTS::run +
    VPRNetEvaluator + DeviceManager + VPRNetModel on libtorch CPU, --devices=cpu,cpu semantics) with all the host threads it can
    use: torchrun's OMP_NUM_THREADS=1 is NOT inherited, libtorch gets one intra-op thread per core, and the arrangement is the
    best of a sweep over actors in {4, cores} (evaluator batch = actors, README.md:87) x evaluator threads in {1, 2}
    (SURVEY.md 8d).  `steps` timed steps after `warmup` untimed ones, each a bounded sample (the same number of searches per
    actor), sized so that the whole run takes about `seconds_target`."""
    drv = os.path.join(ROOT, "oracle", "_ref", "ref_driver")
    cores = os.cpu_count() or 1
    if not os.path.exists(drv):
        return None
    env = {k: v for k, v in os.environ.items() if k not in ("OMP_NUM_THREADS", "MKL_NUM_THREADS", "OMP_PROC_BIND", "GOMP_CPU_AFFINITY")}
    actors = sorted({4, max(4, min(cores, 64))})
    base = [drv, "bench", "game=" + game, "sims=%d" % sims, "devices=2", "torch_threads=%d" % cores]
    out = subprocess.run(base + ["actors=" + ",".join(map(str, actors)), "eval_threads=1,2", "moves=1", "steps=1", "warmup=1"],
                         capture_output=True, text=True, timeout=1200, env=env)
    if out.returncode != 0:
        return dict(error=out.stderr[-300:])
    cal = json.loads(out.stdout.strip().splitlines()[-1])
    per_move = max(cal["seconds"], 1e-3)                       # one timed step of one move per actor, best arrangement
    moves = int(max(1, min(50, round(seconds_target / max(1, steps + warmup) / per_move))))
    out = subprocess.run(base + ["actors=%d" % cal["actors"], "eval_threads=%d" % cal["eval_threads"], "moves=%d" % moves,
                                 "steps=%d" % steps, "warmup=%d" % warmup], capture_output=True, text=True, timeout=3600, env=env)
    if out.returncode != 0:
        return dict(error=out.stderr[-300:])
    best = json.loads(out.stdout.strip().splitlines()[-1])
    best["sweep"] = cal["sweep"]
    best["cores"] = cores
    return best


def reference_sample_text(res, sims):
    return ("%d actor threads x %d moves x %d sims per step, %d timed steps; evaluator batch %d, %d evaluator thread(s) per queue, 2 libtorch-CPU "
            "model replicas, %s torch intra-op threads (best of the sweep %s)") % (
        res["actors"], res["moves"], sims, res["steps"], res["batch"], res["eval_threads"], res["torch_threads"],
        json.dumps([[x["actors"], x["eval_threads"], round(x["sims_per_s"])] for x in res.get("sweep", [])]))


def workload_config(desc, T, sims):
    """`config` of the JSON line: identical in both arms (the reference arm runs a bounded sample of the same workload)."""
    return dict(workload=desc, trees_per_gpu=T, sims_per_move=sims, weights="random-init", data="synthetic observations of the config's shape",
                l2="working set (node pool) exceeds the 126 MB L2; no explicit flush")


_REAL_STDOUT = None


def emit(line):
    import ctypes
    sys.stdout.flush()
    ctypes.CDLL(None).fflush(None)   # C stdio buffers of the libraries, while fd 1 is still stderr
    if _REAL_STDOUT is not None:
        os.dup2(_REAL_STDOUT, 1)
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=None, help="timed steps; default: as many as --seconds of searches take")
    ap.add_argument("--seconds", type=float, default=5.0, help="length of the timed region when --steps is not given, and of the sustained leg")
    ap.add_argument("--no-extras", action="store_true", help="skip the sustained leg and the other workloads' short lines")
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c4d", choices=sorted(WORKLOADS))
    ap.add_argument("--precision", default="auto", choices=["auto", "bf16", "fp32"])
    ap.add_argument("--trees", type=int, default=0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-seconds", type=float, default=None, help="CPU-arm time budget (default: 15 s beside our arm, 60 s for --impl reference)")
    args = ap.parse_args()
    # stdout carries exactly one JSON line.  Libraries print there too (NCCL's "NCCL version ..." banner under torchrun), so
    # file descriptor 1 points at stderr while the work runs and is put back for the line itself (emit()).
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    wl = WORKLOADS[args.workload]
    preset_name = wl.get("preset", args.workload)

    import mzb200
    cfg0 = mzb200.preset(preset_name)
    sims = cfg0.num_simulations

    T = args.trees or wl["trees"]
    if args.impl == "reference":
        if rank != 0:
            return 0
        t0 = time.time()
        K, Wm = args.steps or 5, args.warmup
        res = run_reference_cpu(preset_name, args.cpu_seconds or 60.0, sims, steps=K, warmup=Wm)
        if res is None or "error" in res:
            # the oracle port always exists: time the C/torch-CPU restatement instead
            res = run_port_cpu(preset_name, args.cpu_seconds or 30.0, sims)
            v, kind, step_ms, sample, K = res["sims_per_s"], "port", 1000.0 * res["seconds"], res["sample"], 1
        else:
            v, kind, step_ms = res["sims_per_s"], "reference", 1000.0 * res["seconds"] / K
            sample = reference_sample_text(res, sims)
        line = dict(metric=METRIC, value=v, unit=UNIT, n_gpus=args.gpus, steps=K, warmup=Wm,
                    ms_per_step=step_ms, higher_is_better=True, scaling="weak", vs_baseline=None,
                    dtype="f32", data="synthetic", impl="reference",
                    config=workload_config(wl["desc"], T, sims),
                    cpu_baseline=dict(value=v, unit=UNIT, cores=res.get("cores", os.cpu_count()), kind=kind, sample=sample),
                    e2e=dict(value=v, unit=UNIT, h2d_bytes_per_step=0, d2h_bytes_per_step=0),
                    reference_arm=dict(note="thread-per-actor MCTS::run over VPRNetEvaluator/DeviceManager/VPRNetModel on libtorch CPU; "
                                            "each step is a bounded sample of the workload (trees_per_gpu names the GPU arm's cohort)",
                                       step_seconds=res.get("step_seconds"), sweep=res.get("sweep")),
                    wall_s=time.time() - t0)
        emit(line)
        return 0

    import numpy as np
    import torch
    import torch.distributed as dist
    from mzb200.weights import random_weights, recurrent_flops

    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    precision = {"bf16": mzb200.BF16, "fp32": mzb200.FP32_CHECK}.get(args.precision)
    if precision is None:
        try:
            mzb200.Engine(mzb200.preset("tiny", max_trees=4, precision=mzb200.BF16), local_rank).close()
            precision = mzb200.BF16
        except mzb200.MzbError as e:
            if "not available" not in str(e):
                raise
            precision = mzb200.FP32_CHECK
    pk = peaks()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def measure(workload, T, steps, warmup, seconds, extras):
        """One workload on this rank's GPU: W warm-up + K timed searches with device-resident observations (value), the same
        through host buffers (e2e), and the dominant kernel alone (roofline).  steps=None: as many as `seconds` take."""
        w = WORKLOADS[workload]
        pname = w.get("preset", workload)
        cfg = mzb200.preset(pname, max_trees=T, precision=precision, seed=1000 * rank)
        eng = mzb200.Engine(cfg, local_rank)
        # weights: rank 0 draws them, every other rank receives them with ONE ncclBroadcast of the packed blob
        W = random_weights(cfg, seed=0 if rank == 0 else 1 + rank)
        eng.load_weights(W)
        out = dict()
        if world > 1:
            uid = [mzb200.Engine.nccl_unique_id() if rank == 0 else None]
            dist.broadcast_object_list(uid, src=0)
            eng.nccl_init(uid[0], rank, world)
            torch.cuda.synchronize()
            t0 = time.time()
            eng.broadcast_weights(0)          # first call: includes NCCL's lazy connection set-up
            out["weight_sync_first_ms"] = 1000 * (time.time() - t0)
            dist.barrier()
            t0 = time.time()
            for _ in range(5):                # steady state = what every model_sync_interval costs
                eng.broadcast_weights(0)
            out["weight_sync_ms"] = 1000 * (time.time() - t0) / 5
            out["weight_sync_bytes"] = int(eng.get_shapes().weight_blob_bytes)
            sums = [None] * world
            dist.all_gather_object(sums, eng.weights_checksum())
            assert len(set(sums)) == 1, "weight broadcast left replicas different"
        A, nsim = cfg.num_actions, cfg.num_simulations
        flat = eng.shapes.initial_flat_size
        rng = np.random.default_rng(rank)
        noise = np.ones(T, np.uint8)
        ring = bool(w.get("ring"))
        if ring:
            # reanalyze (self_play.cpp:186-225) as BASELINE config 4 defines it: positions of stored games, each with the legal set
            # and to_play its history recorded, exploration noise on; the frames travel through the device-resident observation ring
            from mzb200 import positions as P
            frames, masks, to_play, n_hist = P.sample_positions(T, seed=rank)
            out["positions"] = "%d positions drawn with replacement from %d stored Connect4 games (tests/golden/selfplay_c4.txt)" % (T, n_hist)
            slots = np.arange(T, dtype=np.int32)
            led_by = np.zeros(T, np.int32)
            eng.ring_create()
            eng.ring_push(slots, frames, led_by)
            obs_host = torch.from_numpy(frames).pin_memory()
        else:
            obs_host = torch.from_numpy(rng.random((T, flat), dtype=np.float32)).pin_memory()
            obs_dev = obs_host.cuda()
            masks = np.full(T, (1 << A) - 1, np.uint32)
            to_play = (np.arange(T) % cfg.num_players).astype(np.int32)
        split = dict(root_ms=0.0, sim_loop_ms=0.0)   # rank-local split of the device time: root (initial inference + expansion) | simulations

        def timed(fn, k):
            barrier()
            dev_ms, t0 = 0.0, time.time()
            launches = 0
            for _ in range(k):
                fn()
                tm = eng.last_search_timing()
                dev_ms += tm["total_ms"]
                split["root_ms"] += tm["root_ms"]
                split["sim_loop_ms"] += tm["sim_loop_ms"]
                launches += tm["kernel_launches"]
            barrier()
            wall_ms = 1000 * (time.time() - t0)
            if world > 1:
                t = torch.tensor([dev_ms, wall_ms], device="cuda", dtype=torch.float64)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                dev_ms, wall_ms = float(t[0]), float(t[1])
            return dev_ms, wall_ms, launches

        if ring:
            def run_host():   # one new frame per position from the host, then the ring-fed search
                eng.ring_reset(slots)
                eng.ring_push(slots, obs_host.numpy(), led_by)
                return eng.run_ring(slots, masks, to_play, noise)
            run_dev = lambda: eng.run_ring(slots, masks, to_play, noise)   # frames already in the ring (HBM)
        else:
            run_dev = lambda: eng.run(None, masks, to_play, noise, device_obs_ptr=obs_dev.data_ptr())
            run_host = lambda: eng.run(obs_host.numpy(), masks, to_play, noise)
        for _ in range(warmup):
            run_dev()
        if steps is None:   # --seconds: derive K from one probe step (all ranks agree on the maximum)
            probe_ms, _, _ = timed(run_dev, 1)
            steps = int(max(3, min(20000, round(1000.0 * seconds / max(probe_ms, 1e-3)))))
            if world > 1:
                t = torch.tensor([steps], device="cuda", dtype=torch.int64)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                steps = int(t[0])
        split["root_ms"] = split["sim_loop_ms"] = 0.0
        sampler = ClockSampler(local_rank)
        sampler.start()
        dev_ms, wall_ms, launches = timed(run_dev, steps)
        clocks = sampler.stop()
        step_split = dict(root_ms_per_step=split["root_ms"] / steps, sim_loop_ms_per_step=split["sim_loop_ms"] / steps,
                          sims_per_s_without_initial_inference=T * nsim * steps / (split["sim_loop_ms"] / 1000.0) * world)
        run_host()
        e2e_steps = steps if steps <= 200 else max(20, steps // 4)
        e2e_dev_ms, e2e_wall_ms, _ = timed(run_host, e2e_steps)
        value = world * T * nsim * steps / (dev_ms / 1000.0)
        e2e_value = world * T * nsim * e2e_steps / (e2e_wall_ms / 1000.0)   # host clock: includes H2D, D2H and host-side noise generation
        # bytes that cross PCIe per step through mzb_search_batch: observations + per-tree inputs + Dirichlet noise (f64) + the
        # slots' mt19937 states BOTH ways (the gamma draws run on the host on the slots' own streams, engine_search.cu) | results
        mt_bytes = T * (624 * 4 + 4)
        h2d = T * flat * 4 + T * (4 + 4 + 1) + T * A * 8 + mt_bytes
        d2h = T * (8 + 4 + 4 + A * (4 + 4 + 8 + 8)) + mt_bytes
        out.update(steps=steps, value=value, ms_per_step=dev_ms / steps, wall_ms_per_step=wall_ms / steps, clocks=clocks,
                   gpu_launches=int(launches), step_split=step_split, sims=nsim, T=T, cfg=cfg, W=W,
                   e2e=dict(value=e2e_value, unit=UNIT, h2d_bytes_per_step=int(h2d), d2h_bytes_per_step=int(d2h),
                            ms_per_step=e2e_wall_ms / e2e_steps, device_ms_per_step=e2e_dev_ms / e2e_steps, steps=e2e_steps,
                            rng_state_round_trip_bytes_each_way=int(mt_bytes)),
                   node_pool_gb=eng.get_shapes().node_pool_bytes / 1e9, recurrent_flat_size=eng.shapes.recurrent_flat_size)

        # roofline of the dominant kernel: the residual towers of recurrent_inference (tensor bound), timed ALONE with CUDA events
        # over back-to-back launches whose inputs rotate through the whole node pool (> L2).  50 launches (~12 ms at full clocks)
        # are a burst measurement -> burst peak; the sustained leg below runs for seconds -> sustained peak.
        frec = recurrent_flops(cfg)
        rec_passes = 50
        rec_ms, rec_launches = eng.bench_recurrent(T, rec_passes)
        states_per_s = T / (rec_ms / 1000.0)
        achieved_tf = frec * states_per_s / 1e12
        traffic = ncu_traffic(workload, T)
        roofline = dict(bound="tensor", achieved=achieved_tf, peak=pk["tf_burst"], unit="TFLOP/s", frac=achieved_tf / pk["tf_burst"],
                        traffic=traffic, traffic_source="ncu --set full capture committed under profiles/ (tower_fused_traffic.json)" if traffic else None,
                        peak_source=pk["source"] + " (burst bf16: %d back-to-back launches, %.1f ms, timed alone)" % (rec_passes, rec_ms * rec_passes),
                        frac_of_sustained_peak=achieved_tf / pk["tf_sustained"],
                        kernel="tower_fused_kernel (one launch = recurrent_inference of all trees)" if rec_launches == 1 else "recurrent_inference towers",
                        flops_per_state=frec, flops_per_launch=frec * T, launch_ms=rec_ms, states_per_s=states_per_s,
                        launches_per_pass=rec_launches)
        if extras:
            # sustained: seconds of back-to-back tower launches and seconds of searches, with the clocks seen
            sp = ClockSampler(local_rank)
            sp.start()
            n_pass = int(max(50, min(200000, round(1000.0 * seconds / max(rec_ms, 1e-3)))))
            s_ms, _ = eng.bench_recurrent(T, n_pass)
            ck = sp.stop()
            s_tf = frec * T / (s_ms / 1000.0) / 1e12
            roofline["sustained"] = dict(achieved=s_tf, peak=pk["tf_sustained"], frac=s_tf / pk["tf_sustained"], launches=n_pass,
                                         seconds=s_ms * n_pass / 1000.0, launch_ms=s_ms, clocks=ck)
            if dev_ms < 1000.0 * seconds * 0.8:
                k2 = int(max(steps, round(1000.0 * seconds / (dev_ms / steps))))
                sp = ClockSampler(local_rank)
                sp.start()
                d2, w2, _ = timed(run_dev, k2)
                out["sustained"] = dict(value=world * T * nsim * k2 / (d2 / 1000.0), steps=k2, seconds=w2 / 1000.0, ms_per_step=d2 / k2, clocks=sp.stop())
        if world > 1:
            # the hand-over DURING self-play (learner.cpp:93-101 while actors search): rank 0 gets new weights, every rank receives
            # them into its shadow blob on the side stream while a search runs, and installs them between two searches
            before = eng.weights_checksum()
            if rank == 0:
                eng.load_weights(random_weights(cfg, seed=4321))
            dist.barrier()
            run_dev()
            plain_ms, _, _ = timed(run_dev, 3)
            t0 = time.time()
            eng.broadcast_weights_begin(0, rank)
            begin_host_ms = 1000 * (time.time() - t0)
            over_ms, _, _ = timed(run_dev, 3)          # searches on the OLD weights while the broadcast is in flight
            t0 = time.time()
            eng.weights_commit()
            run_dev()                                  # first search on the new weights
            commit_ms = 1000 * (time.time() - t0)
            sums = [None] * world
            dist.all_gather_object(sums, eng.weights_checksum())
            assert len(set(sums)) == 1, "overlapped weight hand-over left replicas different"
            assert rank == 0 or sums[rank] != before, "overlapped weight hand-over did not install the new weights"
            out["weight_sync_overlapped"] = dict(begin_call_ms=begin_host_ms, search_ms_per_step_plain=plain_ms / 3,
                                                 search_ms_per_step_during_broadcast=over_ms / 3, commit_plus_one_search_ms=commit_ms)
            t = torch.tensor([states_per_s], device="cuda", dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
            roofline["states_per_s_all_gpus"] = float(t[0])
        out["roofline"] = roofline
        out["states_per_s"] = states_per_s
        eng.close()
        return out

    extras = not args.no_extras
    m = measure(args.workload, T, args.steps, args.warmup, args.seconds, extras)
    cfg, W, A = m["cfg"], m["W"], m["cfg"].num_actions
    line = dict(metric=METRIC, value=m["value"], unit=UNIT, n_gpus=world, steps=m["steps"], warmup=args.warmup,
                ms_per_step=m["ms_per_step"], higher_is_better=True, scaling="weak", vs_baseline=None,
                dtype="bf16" if precision == mzb200.BF16 else "f32", data="synthetic",
                config=workload_config(wl["desc"], T, sims),
                timing="CUDA events around each search on the engine's stream; max over ranks; node pool %.2f GB per GPU" % m["node_pool_gb"],
                wall_ms_per_step=m["wall_ms_per_step"], clocks=m["clocks"], e2e=m["e2e"],
                gpu_launches=m["gpu_launches"], roofline=m["roofline"], recurrent_states_per_s=m["states_per_s"], step_split=m["step_split"])
    for k in ("weight_sync_ms", "weight_sync_first_ms", "weight_sync_bytes", "weight_sync_overlapped", "sustained"):
        if k in m:
            line[k] = m[k]

    if rank == 0 and world == 1:
        # tree kernels alone (HBM bound): same search with the network stubbed by the SE evaluator, which still moves one hidden
        # state parent -> child per simulation.  At the workload's cohort the step is launch/latency bound (1024 warps); the
        # second figure uses a cohort large enough for HBM to be the limit.
        S_h = 4 * m["recurrent_flat_size"]
        d = 4
        b_sim = 2 * S_h + d * 28 * A + 32 * A + 36 * (d + 1) + (4 + 16 + 4 * A)

        def tree_leg(Tt):
            scfg = mzb200.preset(preset_name, max_trees=Tt, evaluator=mzb200.EVAL_SYNTHETIC)
            seng = mzb200.Engine(scfg, local_rank)
            keys = (np.arange(Tt) + 1).astype(np.float32)
            mk = np.full(Tt, (1 << A) - 1, np.uint32)
            tp = (np.arange(Tt) % cfg.num_players).astype(np.int32)
            nz = np.ones(Tt, np.uint8)
            for _ in range(3):
                seng.run(keys, mk, tp, nz)
            ms = 0.0
            for _ in range(5):
                seng.run(keys, mk, tp, nz)
                ms += seng.last_search_timing()["sim_loop_ms"]
            seng.close()
            sims_s = 5 * Tt * sims / (ms / 1000.0)
            return dict(bound="hbm", achieved=b_sim * sims_s / 1e9, peak=pk["hbm"], unit="GB/s", frac=b_sim * sims_s / 1e9 / pk["hbm"],
                        traffic=None, bytes_per_sim=b_sim, sims_per_s=sims_s, trees=Tt, us_per_simulation_step=ms * 1000.0 / (5 * sims),
                        kernel="tree_step + SE stub (outputs + bulk-copy hidden-state move through shared memory)")

        # BASELINE metric M2: recurrent_inference states/s with device-resident inputs, by batch size
        frec = m["roofline"]["flops_per_state"]
        by_batch = {}
        for B in (1024, 4096, 16384):
            pool_bytes = B * (sims + 3) * m["recurrent_flat_size"] * 2
            if pool_bytes > 60e9:
                continue
            beng = mzb200.Engine(mzb200.preset(preset_name, max_trees=B, precision=precision), local_rank)
            beng.load_weights(W)
            b_ms, _ = beng.bench_recurrent(B, 20)
            by_batch[str(B)] = dict(states_per_s=B / (b_ms / 1000.0), ms_per_pass=b_ms, tflops=frec * B / (b_ms / 1000.0) / 1e12,
                                    frac_of_burst_peak=frec * B / (b_ms / 1000.0) / 1e12 / pk["tf_burst"])
            beng.close()
        line["recurrent_states_per_s_by_batch"] = by_batch
        line["roofline_tree"] = tree_leg(T)
        T_big = 16384
        if T_big > T and T_big * (sims + 1) * S_h < 40e9:
            line["roofline_tree_large_cohort"] = tree_leg(T_big)
        if extras:
            # the other single-GPU configs of BASELINE.json as short driver-visible lines (3 warm-up + 3 timed steps each)
            others = {}
            for name in sorted(WORKLOADS):
                if name == args.workload:
                    continue
                try:
                    o = measure(name, WORKLOADS[name]["trees"], 3, 3, 0.0, False)
                    others[name] = dict(value=o["value"], e2e=o["e2e"]["value"], unit=UNIT, trees=o["T"], sims_per_move=o["sims"], steps=3, warmup=3,
                                        ms_per_step=o["ms_per_step"], tower_tflops=o["roofline"]["achieved"],
                                        tower_frac_of_burst_peak=o["roofline"]["frac"], tower_launch_ms=o["roofline"]["launch_ms"],
                                        tower_launches_per_pass=o["roofline"]["launches_per_pass"],
                                        root_ms_per_step=o["step_split"]["root_ms_per_step"], traffic=o["roofline"]["traffic"])
                except Exception as e:   # a side line must not take the headline down
                    others[name] = dict(error=str(e)[:200])
            line["workloads"] = others
        if not args.no_cpu_baseline:
            res = run_reference_cpu(preset_name, args.cpu_seconds or 15.0, sims, steps=3, warmup=1)
            if res and "error" not in res:
                line["cpu_baseline"] = dict(value=res["sims_per_s"], unit=UNIT, cores=res["cores"], kind="reference",
                                            sample=reference_sample_text(res, sims))
            else:
                res = run_port_cpu(preset_name, args.cpu_seconds or 15.0, sims)
                line["cpu_baseline"] = dict(value=res["sims_per_s"], unit=UNIT, cores=1, kind="port", sample=res["sample"])
    if rank == 0:
        emit(line)
    if world > 1:
        dist.destroy_process_group()
    return 0


def run_port_cpu(game, seconds_target, sims):
    """Fallback CPU baseline when oracle/_ref is absent: the oracle port (C search + torch-CPU network), 1 thread."""
    import numpy as np
    import mzb200
    from mzb200.weights import random_weights
    from oracle import oracle_py as O
    from oracle import vprnet_oracle as V
    cfg = mzb200.preset(game)
    W = random_weights(cfg, 0)
    net = V.VPRNetOracle(V.preset(game), W)
    ocfg = O.make_config(num_players=cfg.num_players, num_simulations=sims, num_actions=cfg.num_actions,
                         lower=cfg.value_lowerbound, upper=cfg.value_upperbound)
    rng = np.random.default_rng(0)
    c, h, w = cfg.observation_shape
    flat = (c * (cfg.stacked_observations + 1) + cfg.stacked_observations * cfg.action_channels) * h * w
    t0 = time.time()
    moves = 0
    while time.time() - t0 < seconds_target:
        hidden = {}

        def ev(parent, action, new_node):
            if parent < 0:
                o = net.initial_inference(rng.random((1, flat), dtype=np.float32))
            else:
                o = net.recurrent_inference(hidden[parent], [action])
            hidden[new_node] = o["hidden"]
            return float(o["value"][0]), float(o["reward"][0]), [float(x) for x in o["policy"][0]]
        m = O.mcts_create(ocfg, moves)
        O.mcts_run(m, list(range(cfg.num_actions)), 0, True, ev)
        O.mcts_destroy(m)
        moves += 1
    sec = time.time() - t0
    return dict(sims_per_s=moves * sims / sec, seconds=sec, sample="%d moves x %d sims, 1 thread, oracle port" % (moves, sims), cores=1)


if __name__ == "__main__":
    sys.exit(main())
