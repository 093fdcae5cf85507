#!/usr/bin/env python
"""
bench.py — batched SafeLife env-steps/s on B200 (BASELINE.json metric), one JSON line.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

Workload (BASELINE.json configs[1]): append-still 26x26 benchmark levels, 4096 lockstep envs per
GPU, seeded random actions; one "step" = one fused SafeLifeEnv.step (execute_action + advance_board
on board and goals + points/performance reward + exit colours) over one batch of 4096 envs.
To keep the timed region HBM-honest the bench cycles through a RING of independent 4096-env
batches whose total state is > 2x the 126 MB L2, so every step reads state that has been evicted.

N > 1 (torchrun): envs shard by index, 4096 per GPU (weak scaling); the only collective is the
NCCL all-reduce of the statistics vector at the end of the timed region.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

ENVS_PER_GPU = 4096
H = W = 26
BYTES_PER_ENV_STEP = 8 * H * W + 32  # 4 B per cell-update, board + goals, + scalars
L2_BYTES = 126 * 1024 * 1024


def load_traffic(n):
    """DRAM bytes per launch of the step kernel from the committed ncu capture (only valid for the
    configuration it was captured on: 4096 envs)."""
    p = os.path.join(ROOT, "profiles", "r01", "traffic.json")
    if n != ENVS_PER_GPU or not os.path.exists(p):
        return None, None
    with open(p) as f:
        t = json.load(f)
    return t["dram_bytes_read_per_launch"] + t["dram_bytes_write_per_launch"], t["note"]


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    return 6650.0, "fallback"


class ClockSampler(threading.Thread):
    """Samples nvidia-smi clocks / throttle reasons during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.stop_flag = index, [], threading.Event()

    def run(self):
        while not self.stop_flag.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True,
                                     timeout=5).stdout.strip()
                if out:
                    self.samples.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            self.stop_flag.wait(0.1)

    def summary(self):
        self.stop_flag.set()
        self.join(timeout=6)
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unavailable"]}
        sm = sorted(float(s[0]) for s in self.samples)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for k, n in enumerate(names) if any(s[2 + k].lower().startswith("active") for s in self.samples)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(self.samples[0][1]), "reasons": reasons,
                "samples": len(sm)}


def append_still_levels():
    from tests import golden_util as gu
    return gu.levels("append-still")


# ---------------------------------------------------------------------------------------------
# CPU arm: the reference's advance_board.c (oracle/_ref) + C port of the Python rule code
# ---------------------------------------------------------------------------------------------
def cpu_arm(n_envs, steps, warmup, threads):
    from oracle import slo

    lv = append_still_levels()
    kind = "port"
    if slo.have_ref():
        slo.use_reference_advance(True)
        kind = "reference"  # advance_board is the reference's own C (oracle/_ref); the rule code around it is the port
    env = slo.OracleBatchEnv(slo.OraclePool(**lv), np.arange(n_envs) % 100, seed=0, threads=threads)
    rng = np.random.RandomState(2)
    acts = rng.randint(9, size=(steps + warmup, n_envs))
    for t in range(warmup):
        env.step(acts[t])
    t0 = time.perf_counter()
    for t in range(warmup, warmup + steps):
        env.step(acts[t])
    dt = time.perf_counter() - t0
    slo.use_reference_advance(False)
    physics = "reference advance_board.c (oracle/_ref)" if slo.have_ref() else "oracle port"
    return n_envs * steps / dt, dt, kind, physics


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = len(os.sched_getaffinity(0))
    n_envs = ENVS_PER_GPU  # the same 4096-env batch; each step is one env-step of it
    value, dt, kind, physics = cpu_arm(n_envs, args.steps, args.warmup, threads)
    sample = "%d steps of %d append-still envs; physics = %s, rules = C port of safelife_game.py; %d threads" % (
        args.steps, n_envs, physics, threads)
    line = {
        "impl": "reference", "metric": "env_steps_per_s", "value": value, "unit": "env-steps/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u16", "data": "synthetic",
        "config": {"workload": "append-still 26x26, 4096 lockstep envs, random actions, full env step"},
        "cell_updates_per_s": value * 2 * H * W,
        "cpu_baseline": {"value": value, "unit": "env-steps/s", "cores": threads, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": "env-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))


# ---------------------------------------------------------------------------------------------
# B200 arm
# ---------------------------------------------------------------------------------------------
def run_b200(args):
    import torch

    from safelife_k2_b200 import _lib, distributed
    from safelife_k2_b200.safelife_env import SafeLifeEnv
    from safelife_k2_b200.safelife_game import LevelPool

    # NCCL may print its version banner on stdout when the communicator is created (first collective):
    # keep stdout for the one JSON line by pointing fd 1 at stderr until that has happened
    sys.stdout.flush()
    saved_stdout = os.dup(1)
    os.dup2(2, 1)
    try:
        rank, world, local = distributed.init_from_env()
        torch.cuda.set_device(local)
        dev = torch.device("cuda", local)
        if world > 1:
            torch.distributed.barrier(device_ids=[local])
            distributed.max_over_ranks(0.0, dev)
            torch.cuda.synchronize()
    finally:
        sys.stdout.flush()
        os.dup2(saved_stdout, 1)
        os.close(saved_stdout)
    n = args.envs
    lo = rank * n
    lv = append_still_levels()
    pool = LevelPool(lv, device=dev)
    state_bytes = n * (2 * 64 * H + 64)  # resident: 64*H B per board, 64 B scalars
    ring = max(4, int(np.ceil(2.2 * L2_BYTES / state_bytes)))
    K, Wm = args.steps, max(3, args.warmup)

    envs = []
    for r in range(ring):
        e = SafeLifeEnv(pool, num_envs=n, env_id0=lo + r * n * world, global_counter=None,
                        compute_obs=False, auto_reset=True)
        e.seed(0)
        e.reset()
        envs.append(e)
    rng = np.random.RandomState(2 + rank)
    n_act = 64
    acts = torch.from_numpy(rng.randint(9, size=(n_act, n)).astype(np.uint8)).to(dev)
    stats = torch.zeros(8, dtype=torch.int64, device=dev)
    lib = _lib.load()
    stream = torch.cuda.current_stream()

    def step(i):
        e = envs[i % ring]
        g = e.game
        _lib.check(lib.slb_env_step(pool.c, n, g.board_res.data_ptr(), g.goals_res.data_ptr(), g.state.data_ptr(),
                                    acts[i % n_act].data_ptr(), 1000, 0, g.env_id0, e._reward.data_ptr(),
                                    e._done.data_ptr(), e._stride(), stream.cuda_stream))

    for i in range(Wm):
        step(i)
    torch.cuda.synchronize()
    # One pass over the ring (ring consecutive env-step launches, one per batch) is captured as a
    # CUDA graph so that the host's launch rate (Python + ctypes) is not what gets measured; the
    # remainder K % ring is launched directly.  Every timed step is still its own kernel launch.
    graph = None
    if not args.no_graph and K >= ring:
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            stream = side
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, stream=side):
                for i in range(Wm, Wm + ring):
                    step(i)
        stream = torch.cuda.current_stream()
        torch.cuda.synchronize()
        graph.replay()  # the capture itself does not execute: run the ring once as extra warm-up
        torch.cuda.synchronize()
    if world > 1:
        torch.distributed.barrier()
    sampler = ClockSampler(local)
    sampler.start()
    l0 = _lib.launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    ev0.record(stream)
    n_replays = K // ring if graph is not None else 0
    for _ in range(n_replays):
        graph.replay()
    for i in range(Wm + n_replays * ring, Wm + K):
        step(i)
    e_last = envs[(Wm + K - 1) % ring]
    _lib.check(lib.slb_accumulate_stats(n, e_last.game.state.data_ptr(), e_last._reward.data_ptr(),
                                        e_last._done.data_ptr(), stats.data_ptr(), stream.cuda_stream))
    ev1.record(stream)
    torch.cuda.synchronize()
    launches = _lib.launch_count() - l0 + n_replays * ring  # graph replays launch `ring` kernels each
    ms = ev0.elapsed_time(ev1)
    distributed.all_reduce_stats(stats)
    if world > 1:
        torch.distributed.barrier()
    ms = distributed.max_over_ranks(ms, dev)
    clocks = sampler.summary()

    # ---- informational: the same ring with even / odd batches on two streams (fork-join graph), so that
    # one batch's load / drain overlaps the other's compute.  Not the headline: `value` stays single-stream.
    overlap = None
    if graph is not None:
        s2 = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            graph2 = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph2, stream=side):
                s2.wait_stream(side)
                for i in range(ring):
                    stream = side if i % 2 == 0 else s2
                    step(i)
                side.wait_stream(s2)
        stream = torch.cuda.current_stream()
        torch.cuda.synchronize()
        graph2.replay()
        torch.cuda.synchronize()
        reps = max(1, min(n_replays, 2000))
        o0, o1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        o0.record(stream)
        for _ in range(reps):
            graph2.replay()
        o1.record(stream)
        torch.cuda.synchronize()
        oms = distributed.max_over_ranks(o0.elapsed_time(o1), dev) / (reps * ring)
        overlap = {"streams": 2, "value": n * world / (oms * 1e-3), "ms_per_step": oms,
                   "what": "same ring, even / odd batches on two streams inside one graph"}

    # ---- e2e: host-buffer steps through the public API (SafeLifeEnv.host_stepper -> slb_stepper_step / _wait) ----
    # Every step moves that step's actions from pinned host memory to the device (in the launch parameters),
    # runs the fused step and brings reward + done back to pinned host memory (written by the step's last
    # CTA).  `depth` independent env batches are kept in flight on as many streams (the pipelined actor
    # loop: while one batch's results travel back, the next batches step); each step's outputs are waited
    # for before that batch's buffers are reused.
    e2e_steps = min(K, 5000)
    depth = int(os.environ.get("SLB_BENCH_DEPTH", "6"))
    h_act = torch.from_numpy(rng.randint(9, size=(n_act, n)).astype(np.uint8)).pin_memory()
    h_out = [torch.zeros(5 * n, dtype=torch.uint8).pin_memory() for _ in range(depth)]
    streams = [torch.cuda.Stream(device=dev) for _ in range(depth)]
    a_ptr = [h_act[k].data_ptr() for k in range(n_act)]
    o_ptr = [(o.data_ptr(), o.data_ptr() + 4 * n) for o in h_out]
    torch.cuda.synchronize()

    # the public hot-loop API: one bound stepper per (env batch, stream) — SafeLifeEnv.host_stepper
    hs = [[e.host_stepper(st) for st in streams] for e in envs]

    def e2e_loop(count):
        for i in range(count):
            s = i % depth
            if i >= depth:
                hs[(i - depth) % ring][s].wait()  # outputs of step i - depth are now in h_out[s]
            hs[i % ring][s].step(a_ptr[i % n_act], o_ptr[s][0], o_ptr[s][1])
        for s in range(depth):
            hs[0][s].wait()

    e2e_warm = max(2 * ring + depth, 4000)  # every batch's stepper exists and the launch path has settled
    e2e_loop(e2e_warm)
    if world > 1:
        torch.distributed.barrier()
    if os.environ.get("SLB_BENCH_DEBUG"):
        for _ in range(6):
            t0 = time.perf_counter()
            e2e_loop(500)
            print("e2e chunk of 500: %.1f us/step" % ((time.perf_counter() - t0) / 500 * 1e6), file=sys.stderr)
    t0 = time.perf_counter()
    e2e_loop(e2e_steps)
    e2e_s = distributed.max_over_ranks(time.perf_counter() - t0, dev)
    # the same, strictly one step at a time (launch, wait, read), for reference
    t0 = time.perf_counter()
    for i in range(500):
        hs[i % ring][0].step(a_ptr[i % n_act], o_ptr[0][0], o_ptr[0][1])
        hs[i % ring][0].wait()
    e2e_serial_s = distributed.max_over_ranks(time.perf_counter() - t0, dev) / 500

    if rank != 0:
        return
    value = n * world * K / (ms * 1e-3)
    peak, peak_src = load_peaks()
    achieved = n * BYTES_PER_ENV_STEP / (ms * 1e-3 / K) / 1e9
    traffic, traffic_note = load_traffic(n)
    line = {
        "metric": "env_steps_per_s", "value": value, "unit": "env-steps/s", "n_gpus": world,
        "steps": K, "warmup": Wm, "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u16", "data": "synthetic",
        "config": {"workload": "append-still 26x26, %d lockstep envs per GPU, random actions, full env step " % n +
                               "(execute_action + advance_board x2 + reward + exit colours), k=1 step per launch",
                   "envs_per_gpu": n, "board": [H, W], "l2": "ring of %d independent batches (%.0f MB state) > L2" % (
                       ring, ring * state_bytes / 1e6)},
        "cell_updates_per_s": value * 2 * H * W,
        "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "traffic": traffic, "traffic_note": traffic_note,
                     "peak_source": peak_src + " (MEASURED_PEAKS.json hbm_gbs)" if peak_src == "measured" else peak_src,
                     "algorithmic_bytes_per_launch": n * BYTES_PER_ENV_STEP,
                     "bytes_per_env_step": BYTES_PER_ENV_STEP, "kernel": "res_env_rollout_kernel, T=1 (per launch: %d envs)" % n},
        "e2e": {"value": n * world * e2e_steps / e2e_s, "unit": "env-steps/s", "h2d_bytes_per_step": n,
                "d2h_bytes_per_step": 5 * n, "steps": e2e_steps,
                "what": "SafeLifeEnv.host_stepper(stream).step / .wait -> slb_stepper_step / _wait: per step the uint8 actions "
                        "are read from pinned host memory and travel to the device in the launch parameters, the fused step "
                        "runs, and its last CTA writes int32 reward + uint8 done to pinned host memory and publishes a "
                        "sequence number the host waits on; %d independent batches in flight on %d streams; %d untimed "
                        "warm-up steps (the driver's launch path settles after ~2 500 of these 4 KB-parameter launches: "
                        "10 -> 6.3 us per step)" % (depth, depth, e2e_warm),
                "serial_value": n * world / e2e_serial_s,
                "serial_what": "the same calls strictly one step at a time: step, wait, read"},
        "stats_allreduce": distributed.stats_dict(stats),
    }
    if overlap is not None:
        line["two_stream_overlap"] = overlap
    # CPU baseline on this box's host cores, bounded sample (rank 0, N=1 only)
    if world == 1 and not args.no_cpu:
        threads = len(os.sched_getaffinity(0))
        v0, dt0, _, _ = cpu_arm(n, 20, 2, threads)          # calibrate, then ~10 s of CPU work
        cpu_steps = int(min(20000, max(50, 10.0 / (dt0 / 20))))
        v, dt, kind, physics = cpu_arm(n, cpu_steps, 2, threads)
        line["cpu_baseline"] = {"value": v, "unit": "env-steps/s", "cores": threads, "kind": kind,
                                "sample": "%d steps of the same %d-env batch (%.1f s of wall clock on %d threads); "
                                          "physics = %s, rules = C port of safelife_game.py" % (cpu_steps, n, dt, threads, physics)}
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=None)
    ap.add_argument("--warmup", type=int, default=None)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--envs", type=int, default=ENVS_PER_GPU, help="envs per GPU per step (default: the BASELINE config)")
    ap.add_argument("--no-graph", action="store_true", help="launch every step from Python (host-bound)")
    args = ap.parse_args()
    if args.impl == "reference":
        args.steps = args.steps if args.steps is not None else 20
        args.warmup = args.warmup if args.warmup is not None else 2
        run_reference(args)
    else:
        args.steps = args.steps if args.steps is not None else 100000
        args.warmup = args.warmup if args.warmup is not None else 50
        run_b200(args)
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized():
            dist.destroy_process_group()


if __name__ == "__main__":
    main()
