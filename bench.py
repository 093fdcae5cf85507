#!/usr/bin/env python
"""
bench.py — batched SafeLife env-steps/s on B200 (BASELINE.json metric), one JSON line.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

Workload (BASELINE.json configs[1]): append-still 26x26 benchmark levels, 4096 lockstep envs per GPU, seeded
random actions; one "step" = one fused SafeLifeEnv.step (execute_action + advance_board on board and goals +
points/performance reward + exit colours) over one batch of 4096 envs = ONE kernel launch.  To keep the timed
region HBM-honest the steps cycle through a RING of independent 4096-env batches whose total state is > 2 x the
126 MB L2, so every step reads state that has been evicted.

How K steps are timed, for any K: the K step launches (W untimed ones first) are captured into ONE CUDA graph —
so the host's Python / ctypes launch rate is not what is measured — and, because K = 20 steps take < 0.1 ms, that
K-step graph is replayed R times back to back inside the timed region until it is >= 200 ms long (R is fixed
before the timed region from a calibration replay).  The line reports `steps: K`, `config.replays: R`,
`ms_per_step` = timed region / (R * K); the statistics kernel and the NCCL all-reduce of the statistics vector
(BASELINE config 5's "with NCCL reward-stat all-reduce") are INSIDE the timed region.

N > 1 (torchrun): envs shard by index, 4096 per GPU (weak scaling, `value`); the line's `strong` block carries
BASELINE config 5: a fixed total (65536 envs @26x26 and 4096 envs @256x256) sharded over the N ranks.
At N = 1 the `configs` block carries BASELINE configs 3 and 4 and the other measured paths.
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
WORKLOAD = ("append-still 26x26, 4096 lockstep envs per GPU, seeded random actions, full env step "
            "(execute_action + advance_board x2 + reward + exit colours), k=1 step per launch")
MIN_TIMED_MS = 200.0


class ClockSampler(threading.Thread):
    """Samples SM clocks / throttle reasons of one GPU while the timed regions run: NVML every 5 ms when pynvml
    is importable (the GPU is found by its UUID, so CUDA_VISIBLE_DEVICES does not matter), else nvidia-smi."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
    BITS = [0x8, 0x40, 0x20, 0x4]  # nvmlClocksEventReason{HwSlowdown, HwThermalSlowdown, SwThermalSlowdown, SwPowerCap}

    def __init__(self, index, uuid=None):
        super().__init__(daemon=True)
        self.index, self.uuid, self.samples, self.stop_flag = index, uuid, [], threading.Event()
        self.sm_max, self.source = None, "nvidia-smi"
        self.handle = None
        try:
            import pynvml
            pynvml.nvmlInit()
            h = None
            if uuid:
                try:
                    h = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + str(uuid)).encode())
                except Exception:
                    h = None
            if h is None:
                h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.sm_max = float(pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM))
            self.handle, self.nv, self.source = h, pynvml, "nvml"
        except Exception:
            self.handle = None

    def _reasons_nvml(self):
        nv = self.nv
        fn = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or nv.nvmlDeviceGetCurrentClocksThrottleReasons
        return int(fn(self.handle))

    def run(self):
        while not self.stop_flag.is_set():
            try:
                if self.handle is not None:
                    sm = float(self.nv.nvmlDeviceGetClockInfo(self.handle, self.nv.NVML_CLOCK_SM))
                    r = self._reasons_nvml()
                    self.samples.append((sm, [bool(r & b) for b in self.BITS]))
                    self.stop_flag.wait(0.005)
                    continue
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True,
                                     timeout=5).stdout.strip()
                if out:
                    f = [x.strip() for x in out.split(",")]
                    self.sm_max = float(f[1])
                    self.samples.append((float(f[0]), [x.lower().startswith("active") for x in f[2:6]]))
            except Exception:
                pass
            self.stop_flag.wait(0.02)

    def summary(self):
        self.stop_flag.set()
        self.join(timeout=6)
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unavailable"]}
        sm = sorted(s[0] for s in self.samples)
        reasons = [n for k, n in enumerate(self.NAMES) if any(s[1][k] for s in self.samples)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_min_mhz": sm[0], "sm_max_mhz": self.sm_max, "reasons": reasons,
                "samples": len(sm), "source": self.source}


def append_still_levels():
    from tests import golden_util as gu
    return gu.levels("append-still")


# ---------------------------------------------------------------------------------------------
# CPU arm: the reference's advance_board.c (oracle/_ref) + C port of the Python rule code
# ---------------------------------------------------------------------------------------------
CPU_KIND_REF = "reference-physics+ported-rules"


def cpu_arm(n_envs, steps, warmup, threads):
    """`steps` env-steps of one n_envs batch on `threads` host threads -> (env-steps/s, seconds, kind, physics)"""
    from oracle import slo

    lv = append_still_levels()
    kind = "port"
    if slo.have_ref():
        slo.use_reference_advance(True)
        kind = CPU_KIND_REF  # advance_board is the reference's own C (oracle/_ref); the rule code around it is the port
    env = slo.OracleBatchEnv(slo.OraclePool(**lv), np.arange(n_envs) % 100, seed=0, threads=threads)
    rng = np.random.RandomState(2)
    acts = rng.randint(9, size=(64, n_envs))
    for t in range(warmup):
        env.step(acts[t % 64])
    t0 = time.perf_counter()
    for t in range(warmup, warmup + steps):
        env.step(acts[t % 64])
    dt = time.perf_counter() - t0
    slo.use_reference_advance(False)
    physics = "reference advance_board.c (oracle/_ref)" if slo.have_ref() else "oracle port"
    return n_envs * steps / dt, dt, kind, physics


def run_reference(args):
    """The reference's CPU implementation of the same env step on the box's host cores.  A `step` of this arm is
    one env-step of the same 4096-env batch; each of the K timed steps is repeated `inner` times (a bounded
    sample: the whole run lasts ~10 s) and ms_per_step is the mean time of one batch step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = len(os.sched_getaffinity(0))
    n_envs = ENVS_PER_GPU
    _, dt0, _, _ = cpu_arm(n_envs, max(4, args.warmup), 1, threads)   # calibrate (also the warm-up)
    per = dt0 / max(4, args.warmup)
    budget = float(os.environ.get("SLB_REF_SECONDS", "10"))  # seconds of CPU work in the timed sample
    inner = int(max(1, min(2000, round(budget / (per * args.steps)))))
    value, dt, kind, physics = cpu_arm(n_envs, args.steps * inner, args.warmup, threads)
    sample = ("%d x %d env-steps of one %d-env append-still batch (%.1f s of wall clock on %d threads); physics = %s, "
              "rules = C port of safelife_game.py" % (args.steps, inner, n_envs, dt, threads, physics))
    line = {
        "impl": "reference", "metric": "env_steps_per_s", "value": value, "unit": "env-steps/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * dt / (args.steps * inner), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u16", "data": "synthetic",
        "config": {"workload": WORKLOAD, "envs_per_gpu": n_envs, "board": [H, W], "replays": inner},
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

    import bench_configs as bc
    from safelife_k2_b200 import _lib, distributed
    from safelife_k2_b200.safelife_game import LevelPool
    from tests import golden_util as gu

    # NCCL may print its version banner on stdout when the communicator is created (first collective):
    # keep stdout for the one JSON line by pointing fd 1 at stderr until the line is printed
    sys.stdout.flush()
    saved_stdout = os.dup(1)
    os.dup2(2, 1)
    rank, world, local = distributed.init_from_env()
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        torch.distributed.barrier(device_ids=[local])
        distributed.max_over_ranks(0.0, dev)
        torch.cuda.synchronize()
    n = args.envs
    lv = append_still_levels()
    pool = LevelPool(lv, device=dev)
    K, Wm = args.steps, max(3, args.warmup)
    lib = _lib.load()

    # ---- the ring of independent batches; batch r of rank k plays the envs (r * world + k) * n + [0, n) ----
    envs, state_bytes = bc.ring_envs(pool, n, env_id0=rank * n, id_stride=n * world, min_ring=4)
    ring = len(envs)
    rng = np.random.RandomState(2 + rank)
    n_act = 64
    acts = torch.from_numpy(rng.randint(9, size=(n_act, n)).astype(np.uint8)).to(dev)
    stats = torch.zeros(8, dtype=torch.int64, device=dev)
    launch = bc.step_launcher(pool, envs, acts, lib)

    # W untimed warm-up steps (launched one by one), then the K-step graph.  A graph of K launches touches K
    # consecutive batches of the ring; with K < ring the graph is made of r = ceil(ring / K) passes of K steps
    # so that a batch is touched again only after (almost) the whole ring went by.
    cur = torch.cuda.current_stream().cuda_stream
    for j in range(Wm):
        launch(j, cur)
    torch.cuda.synchronize()
    passes = 1
    if not args.no_graph:
        while True:
            L = K * passes
            gap = L % ring if L % ring else ring  # batches touched between two touches of the same batch
            if L >= ring and (gap - 1) * state_bytes >= 1.5 * L2_BYTES:
                break
            passes += 1
        graph = bc.capture(lambda j, st: launch(Wm + j, st), K * passes, dev)
    last = envs[(Wm + K * passes - 1) % ring]

    def tail():
        # BASELINE config 5: "with NCCL reward-stat all-reduce" — inside the timed region
        _lib.check(lib.slb_accumulate_stats(n, last.game.state.data_ptr(), last._reward.data_ptr(),
                                            last._done.data_ptr(), stats.data_ptr(),
                                            torch.cuda.current_stream().cuda_stream))
        distributed.all_reduce_stats(stats)

    try:
        uuid = torch.cuda.get_device_properties(dev).uuid
    except Exception:
        uuid = None
    sampler = ClockSampler(local, uuid)
    sampler.start()
    l0 = _lib.launch_count()
    if args.no_graph:
        # every step launched from Python (host-bound; not the judged mode — the line says graph: false)
        if world > 1:
            torch.distributed.barrier()
        torch.cuda.synchronize()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record()
        for j in range(K):
            launch(Wm + j, cur)
        tail()
        ev1.record()
        torch.cuda.synchronize()
        ms = distributed.max_over_ranks(ev0.elapsed_time(ev1), dev)
        replays, launches = 1, K + 1
    else:
        ms, reps = bc.time_graph(graph, MIN_TIMED_MS, dev, world, tail)
        replays = reps * passes            # how many times the K steps ran inside the timed region
        launches = reps * passes * K + 1   # + the statistics kernel
    clocks = sampler.summary()
    ms_per_step = ms / (replays * K)
    fs_static = bc.goals_static_fraction(pool, envs)

    # ---- informational: one batch stepped back to back (a policy in the loop on a single batch: every launch
    # depends on the previous one and the batch's 13.7 MB of state stays in L2) ----
    single = None
    if not args.no_graph:
        one = bc.step_launcher(pool, envs[:1], acts, lib)
        g1 = bc.capture(one, 50, dev)
        ms1, r1 = bc.time_graph(g1, 20.0, dev, world)
        single = {"value": n * world / (ms1 / (r1 * 50) * 1e-3), "ms_per_step": ms1 / (r1 * 50),
                  "what": "ONE batch stepped back to back (dependent launches, state L2-resident)"}

    # ---- informational: the ring with even / odd batches on two streams inside one graph (an actor that
    # alternates two independent batches: one batch's load and drain overlap the other's compute).  Not the
    # headline: `value` stays one stream, every launch ordered behind the previous one. ----
    two = None
    if not args.no_graph and ring >= 4:
        L2n = ring - (ring % 2)
        g2 = bc.capture2(lambda j, st: launch(j, st), L2n, dev)
        ms2, r2 = bc.time_graph(g2, 20.0, dev, world)
        two = {"streams": 2, "value": n * world / (ms2 / (r2 * L2n) * 1e-3), "ms_per_step": ms2 / (r2 * L2n),
               "what": "same ring, even / odd batches on two streams inside one graph"}

    # ---- e2e: host-buffer steps through the public API (SafeLifeEnv.host_stepper -> slb_stepper_step / _wait) ----
    # Every step moves that step's actions from pinned host memory to the device (in the launch parameters),
    # runs the fused step and brings reward + done back to pinned host memory (written by the step's last
    # CTA).  `depth` independent env batches are kept in flight on as many streams (the pipelined actor
    # loop: while one batch's results travel back, the next batches step); each step's outputs are waited
    # for before that batch's buffers are reused.
    e2e_steps = max(20000, min(K, 100000))  # >= 20 000 steps (~0.1 s of wall clock) whatever K is
    depth = int(os.environ.get("SLB_BENCH_DEPTH", "7"))
    h_act = torch.from_numpy(rng.randint(9, size=(n_act, n)).astype(np.uint8)).pin_memory()
    h_out = [torch.zeros(5 * n, dtype=torch.uint8).pin_memory() for _ in range(depth)]
    streams = [torch.cuda.Stream(device=dev) for _ in range(depth)]
    a_ptr = [h_act[k].data_ptr() for k in range(n_act)]
    o_ptr = [(o.data_ptr(), o.data_ptr() + 4 * n) for o in h_out]
    torch.cuda.synchronize()
    hs = [[e.host_stepper(st) for st in streams] for e in envs]
    if n % 16 == 0 and os.environ.get("SLB_BENCH_BIND", "1") != "0":
        for row in hs:  # stream k always delivers into pinned block k: let the step write there directly
            for k, h in enumerate(row):
                h.bind_outputs(o_ptr[k][0], o_ptr[k][1])

    def e2e_loop(count):
        for i in range(count):
            s = i % depth
            if i >= depth:
                hs[(i - depth) % ring][s].wait()  # outputs of step i - depth are now in h_out[s]
            hs[i % ring][s].step(a_ptr[i % n_act], o_ptr[s][0], o_ptr[s][1])
        for s in range(depth):
            hs[0][s].wait()

    # the launch path of these 4 KB-parameter launches settles after ~2 500 launches of a process
    # (profiles/r01/i_e2e_warmup_chunks.txt); a training run lives in the settled regime
    e2e_warm = max(2 * ring + depth, 4000)
    e2e_loop(e2e_warm)
    if world > 1:
        torch.distributed.barrier()
    # five back-to-back segments of e2e_steps steps each, every one timed on its own (max over ranks); the
    # reported figure is the MEDIAN segment: a host loop of 4-5 us steps is at the mercy of the box's other
    # processes, and one 0.1 s sample moved by +-10 % between runs
    seg_s = []
    for _ in range(5):
        t0 = time.perf_counter()
        e2e_loop(e2e_steps)
        seg_s.append(distributed.max_over_ranks(time.perf_counter() - t0, dev))
    e2e_s = sorted(seg_s)[len(seg_s) // 2]
    # the same, strictly one step at a time (launch, wait, read)
    serial_steps = 1000
    t0 = time.perf_counter()
    for i in range(serial_steps):
        hs[i % ring][0].step(a_ptr[i % n_act], o_ptr[0][0], o_ptr[0][1])
        hs[i % ring][0].wait()
    e2e_serial_s = distributed.max_over_ranks(time.perf_counter() - t0, dev) / serial_steps
    for e in envs:
        e._drop_steppers()

    # ---- BASELINE config 5 (strong scaling) on every N; configs 3 / 4 and the other paths at N = 1 ----
    strong, configs = None, None
    if not args.headline_only:
        del hs
        strong = {}
        for wl, total in (("mixed_26", 65536), ("big_256", 4096)):
            torch.cuda.empty_cache()
            strong[wl] = bc.strong_config(wl, total, rank, world, dev)
        if world == 1:
            consts = bc.profile_constants()
            configs = {}
            torch.cuda.empty_cache()
            ip = bc.int_peak_config(dev)
            configs["int_peak"] = ip
            c3 = bc.step_config("prune_spawn_16384", gu.levels("prune-spawn"), 16384, 26, 26, dev=dev, consts=consts)
            if c3.get("alu_warp_inst_per_env_step"):
                c3["alu_pipe_frac"] = c3["alu_warp_inst_per_env_step"] * c3["env_steps_per_s"] / ip["lop3_warp_inst_per_s"]
            configs["prune_spawn_16384"] = c3
            torch.cuda.empty_cache()
            configs["side_effects"] = bc.side_effects_config(16384, dev)
            torch.cuda.empty_cache()
            configs["mixed_65536"] = bc.step_config("mixed_65536", bc.mixed_levels(), 65536, 26, 26, dev=dev, consts=consts)
            torch.cuda.empty_cache()
            configs["big_256"] = bc.big_config(4096, dev=dev)
            torch.cuda.empty_cache()
            for c in bc.advance_board_configs(dev=dev):
                if c.get("alu_warp_inst_per_board"):  # the spawner boards sit on the integer roofline, not the HBM one
                    c["alu_pipe_frac"] = (c["alu_warp_inst_per_board"] * c["boards"] / (c["ms_per_call"] * 1e-3)
                                          / ip["lop3_warp_inst_per_s"])
                configs[c.pop("config")] = c
            configs["advance_board_host"] = bc.advance_board_host_config()
            configs["env_step_api"] = bc.env_step_api_config(dev=dev)
            configs["rollout_k32"] = bc.rollout_config(dev=dev)

    if rank != 0:
        return
    value = n * world / (ms_per_step * 1e-3)
    peak, peak_src = bc.hbm_peak()
    achieved = n * BYTES_PER_ENV_STEP / (ms_per_step * 1e-3) / 1e9
    consts = bc.profile_constants().get("headline_step", {}) if n == ENVS_PER_GPU else {}
    traffic = consts.get("dram_bytes_per_launch")
    line = {
        "metric": "env_steps_per_s", "value": value, "unit": "env-steps/s", "n_gpus": world,
        "steps": K, "warmup": Wm, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u16", "data": "synthetic",
        "config": {"workload": WORKLOAD, "envs_per_gpu": n, "board": [H, W],
                   "graph": not args.no_graph, "replays": replays, "timed_ms": ms, "timed_steps": replays * K,
                   "l2": "ring of %d independent batches (%.0f MB state) > L2; steps cycle through it" % (
                       ring, ring * state_bytes / 1e6),
                   "allreduce_in_timed_region": True},
        "cell_updates_per_s": value * 2 * H * W,
        "cell_updates_executed_per_s": value * (2 - fs_static) * H * W,
        "goals_static_fraction": fs_static,
        "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak,
                     "traffic": traffic,
                     "frac_traffic": (traffic / (ms_per_step * 1e-3) / 1e9 / peak) if traffic else None,
                     "traffic_source": consts.get("source"),
                     # the capture's DRAM bytes + the launch's stores (dirty in L2 when the profiled kernel ends)
                     "traffic_with_writeback": (traffic + consts.get("store_bytes_per_launch", 0)) if traffic else None,
                     "frac_traffic_with_writeback": ((traffic + consts.get("store_bytes_per_launch", 0))
                                                     / (ms_per_step * 1e-3) / 1e9 / peak) if traffic else None,
                     "peak_source": peak_src,
                     "algorithmic_bytes_per_launch": n * BYTES_PER_ENV_STEP,
                     "bytes_per_env_step": BYTES_PER_ENV_STEP,
                     "kernel": "res_env_rollout_kernel, T=1 (per launch: %d envs)" % n},
        "e2e": {"value": n * world * e2e_steps / e2e_s, "unit": "env-steps/s", "h2d_bytes_per_step": n,
                "d2h_bytes_per_step": 5 * n, "steps": e2e_steps, "warmup_steps": e2e_warm, "depth": depth,
                "seconds": e2e_s, "segments": len(seg_s),
                "segment_values": [n * world * e2e_steps / x for x in seg_s],
                "what": "SafeLifeEnv.host_stepper(stream).step / .wait -> slb_stepper_step / _wait: per step the uint8 "
                        "actions are read from pinned host memory and travel to the device in the launch parameters, "
                        "the fused step runs, and its last CTA writes int32 reward + uint8 done to pinned host memory "
                        "and publishes a sequence number the host waits on; `depth` independent batches in flight on "
                        "as many streams, wall clock over `steps` steps after `warmup_steps` untimed ones; median of `segments` "
                        "consecutive samples of `steps` steps each (`segment_values`)",
                "serial_value": n * world / e2e_serial_s, "serial_steps": serial_steps,
                "serial_what": "the same calls strictly one step at a time: step, wait, read"},
        "stats_allreduce": distributed.stats_dict(stats),
    }
    if single is not None:
        line["single_batch"] = single
    if two is not None:
        line["two_stream_overlap"] = two
    if strong is not None:
        line["strong"] = strong
    if configs is not None:
        line["configs"] = configs
    # CPU baseline on this box's host cores, bounded sample (rank 0, N=1 only)
    if world == 1 and not args.no_cpu:
        threads = len(os.sched_getaffinity(0))
        v0, dt0, _, _ = cpu_arm(n, 20, 2, threads)          # calibrate, then ~10 s of CPU work
        cpu_steps = int(min(20000, max(50, 10.0 / (dt0 / 20))))
        v, dt, kind, physics = cpu_arm(n, cpu_steps, 2, threads)
        line["cpu_baseline"] = {"value": v, "unit": "env-steps/s", "cores": threads, "kind": kind,
                                "sample": "%d steps of the same %d-env batch (%.1f s of wall clock on %d threads); "
                                          "physics = %s, rules = C port of safelife_game.py" % (cpu_steps, n, dt, threads, physics)}
    sys.stdout.flush()
    os.dup2(saved_stdout, 1)
    os.close(saved_stdout)
    print(json.dumps(line))
    sys.stdout.flush()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=None)
    ap.add_argument("--warmup", type=int, default=None)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--envs", type=int, default=ENVS_PER_GPU, help="envs per GPU per step (default: the BASELINE config)")
    ap.add_argument("--no-graph", action="store_true", help="launch every step from Python (host-bound)")
    ap.add_argument("--headline-only", action="store_true", help="skip the `strong` and `configs` blocks")
    args = ap.parse_args()
    if args.impl == "reference":
        args.steps = args.steps if args.steps is not None else 20
        args.warmup = args.warmup if args.warmup is not None else 3
        run_reference(args)
    else:
        args.steps = args.steps if args.steps is not None else 100
        args.warmup = args.warmup if args.warmup is not None else 10
        if args.steps < 1:
            ap.error("--steps must be >= 1")
        run_b200(args)
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized():
            dist.destroy_process_group()


if __name__ == "__main__":
    main()
