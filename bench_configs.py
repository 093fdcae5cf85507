#!/usr/bin/env python
"""
bench_configs.py — the workloads of BASELINE.json beside the headline (config 2), as functions that
bench.py calls for the `configs` / `strong` blocks of its JSON line, and as a CLI:

    python bench_configs.py [names...] [--quick]

  prune_spawn_16384   config 3: prune-spawn 26x26, 16384 envs, stochastic spawners, full env step, k=1
  side_effects        config 4: side_effect_score counts (1000 samples) of config 3's 16384 final states
  mixed_65536         config 5a: 65536 envs cycled over all 8 v1.0 archives, k=1
  big_256             config 5b: 4096 envs at 256x256 (synthetic levels of SURVEY §8(d)), k=1
  rollout_k32         config 2 with 32 env steps fused per launch (pre-staged actions)
  step_obs / obs      the policy-in-the-loop pair; get_obs 15x15x15 for 4096 envs
  advance_board       the speedups.advance_board contract: uint16 boards in / out, k = 1 and k = 32
  advance_board_host  speedups.advance_board(ndarray): latency of the literal drop-in call
  env_step_api        wall clock of SafeLifeEnv.step (device actions in, observation out), default kwargs
  int_peak            LOP3 / LOP3+IMAD issue-rate microbenchmark (denominator of the integer roofline)

Every device-timed number is taken on a CUDA graph of the launches (replayed until the timed region is
>= min_ms), with CUDA events on the replaying stream, after warm-up; the state a launch touches was last
touched more than one L2 (126 MB) of other traffic ago (rings of independent batches).
"""
import json
import math
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
from safelife_k2_b200 import _lib, distributed  # noqa: E402
from safelife_k2_b200.safelife_env import SafeLifeEnv  # noqa: E402
from safelife_k2_b200.safelife_game import LevelPool  # noqa: E402
from safelife_k2_b200 import side_effects as se  # noqa: E402
from tests import golden_util as gu  # noqa: E402

L2_BYTES = 126 * 1024 * 1024


def hbm_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def profile_constants():
    """Per-kernel constants that only a profiler can give (DRAM bytes per launch, warp instructions per
    env-step and their ALU-pipe share), read from the committed ncu summaries of this round."""
    p = os.path.join(ROOT, "profiles", "r02", "kernel_constants.json")
    if not os.path.exists(p):
        p = os.path.join(ROOT, "profiles", "r01", "kernel_constants.json")
    if os.path.exists(p):
        with open(p) as f:
            return json.load(f)
    return {}


# ---------------------------------------------------------------------------------------------
# timing: a CUDA graph of L launches, replayed
# ---------------------------------------------------------------------------------------------
def capture(launch, L, dev):
    """launch(j, cuda_stream_ptr) for j in 0..L-1, captured on a side stream."""
    side = torch.cuda.Stream(device=dev)
    side.wait_stream(torch.cuda.current_stream())
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.stream(side):
        with torch.cuda.graph(graph, stream=side):
            for j in range(L):
                launch(j, side.cuda_stream)
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    return graph


def capture2(launch, L, dev):
    """launch(j, cuda_stream_ptr) for j in 0..L-1 with even / odd j on two streams (fork / join inside one graph):
    the launches must touch L different batches."""
    side, s2 = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)
    side.wait_stream(torch.cuda.current_stream())
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.stream(side):
        with torch.cuda.graph(graph, stream=side):
            s2.wait_stream(side)
            for j in range(L):
                launch(j, (side if j % 2 == 0 else s2).cuda_stream)
            side.wait_stream(s2)
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    return graph


def time_graph(graph, min_ms=50.0, dev=None, world=1, tail=None, max_reps=200000):
    """Replays `graph` on the current stream until >= min_ms have passed (the repetition count is fixed before
    the timed region from one calibration replay; with several ranks every rank uses the largest count).
    `tail()` runs inside the timed region after the last replay (statistics kernel + all-reduce).
    -> (ms_total, reps): CUDA events on the replaying stream, max over ranks."""
    graph.replay()  # the capture did not execute anything
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    graph.replay()
    e1.record()
    torch.cuda.synchronize()
    # (+ 15 %: the calibration replay runs a little slower than the steady state)
    reps = int(min(max_reps, max(1, math.ceil(1.15 * min_ms / max(e0.elapsed_time(e1), 1e-4)))))
    if world > 1:
        reps = int(distributed.max_over_ranks(float(reps), dev))
        torch.distributed.barrier()
    torch.cuda.synchronize()
    e0.record()
    for _ in range(reps):
        graph.replay()
    if tail is not None:
        tail()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    if world > 1:
        torch.distributed.barrier()
        ms = distributed.max_over_ranks(ms, dev)
    return ms, reps


def ring_envs(pool, n, env_id0=0, id_stride=None, min_bytes=2.2 * L2_BYTES, min_ring=1, **kw):
    """Independent n-env batches whose resident state together exceeds min_bytes.  Batch r plays the envs
    env_id0 + r * id_stride + [0, n)."""
    per = n * (2 * pool.resident_bytes + 80)
    ring = max(min_ring, int(np.ceil(min_bytes / per)))
    id_stride = n if id_stride is None else id_stride
    envs = []
    for r in range(ring):
        e = SafeLifeEnv(pool, num_envs=n, env_id0=env_id0 + r * id_stride, global_counter=None, compute_obs=False,
                        auto_reset=True, **kw)
        e.seed(0)
        e.reset()
        envs.append(e)
    return envs, per


def step_launcher(pool, envs, acts, lib=None):
    lib = lib or _lib.load()
    n, n_act, ring = envs[0].num_envs, acts.shape[0], len(envs)
    args = []
    for e in envs:
        g = e.game
        args.append((g.board_res.data_ptr(), g.goals_res.data_ptr(), g.state.data_ptr(), g.env_id0,
                     e._reward.data_ptr(), e._done.data_ptr(), e._stride()))
    a_ptr = [acts[k].data_ptr() for k in range(n_act)]

    def launch(j, st):
        b, g, s, id0, r, d, stride = args[j % ring]
        _lib.check(lib.slb_env_step(pool.c, n, b, g, s, a_ptr[j % n_act], 1000, 0, id0, r, d, stride, st))

    return launch


def goals_static_fraction(pool, envs):
    """fraction of the envs whose level's goals board is static (its advance_board is skipped, exactly)"""
    flags = pool.level_consts.view(torch.int32)[:, 2].cpu().numpy() & 1
    lv = torch.cat([e.game.level_id for e in envs]).cpu().numpy()
    return float(flags[lv].mean())


def step_config(name, lv, n, h, w, min_ms=50.0, dev=None, consts=None):
    dev = dev or torch.device("cuda", torch.cuda.current_device())
    pool = LevelPool(lv, device=dev)
    envs, per = ring_envs(pool, n)
    acts = torch.from_numpy(np.random.RandomState(3).randint(9, size=(16, n)).astype(np.uint8)).to(dev)
    launch = step_launcher(pool, envs, acts)
    for j in range(len(envs)):
        launch(j, torch.cuda.current_stream().cuda_stream)
    graph = capture(launch, len(envs), dev)
    ms, reps = time_graph(graph, min_ms, dev)
    ms_step = ms / (reps * len(envs))
    hbm, _ = hbm_peak()
    bytes_alg = n * (8 * h * w + 32)
    fs = goals_static_fraction(pool, envs)
    out = {"config": name, "envs": n, "board": [h, w], "ms_per_step": ms_step, "env_steps_per_s": n / ms_step * 1e3,
           "cell_updates_per_s": n * 2 * h * w / ms_step * 1e3,
           "cell_updates_executed_per_s": n * (2 - fs) * h * w / ms_step * 1e3,
           "hbm_frac_algorithmic": bytes_alg / (ms_step * 1e-3) / 1e9 / hbm,
           "ring": len(envs), "timed_steps": reps * len(envs), "timed_ms": ms}
    k = (consts or {}).get(name)
    if k:
        out["warp_inst_per_env_step"] = k.get("warp_inst_per_env_step")
        out["alu_warp_inst_per_env_step"] = k.get("alu_warp_inst_per_env_step")
    return out


def synth_256(rng5, rng6, p):
    vals = np.array([0, 1033, 16, 1041, 521, 32788, 1688], np.uint16)
    pr = np.array([60, 25, 5, 4, 3, 2, 1], float) / 100
    board = vals[rng5.choice(len(vals), size=(p, 256, 256), p=pr)]
    board[:, 127:130, 127:130] = 0
    board[:, 128, 128] = 122
    board[:, 0, 0] = 784
    g = vals[rng6.choice(len(vals), size=(p, 256, 256), p=pr)]
    goals = np.where(g == 1033, 2057, np.where(g == 16, 3584, 0)).astype(np.uint16)
    return {"board": board, "goals": goals, "agent_loc": np.tile([128, 128], (p, 1)), "orientation": np.ones(p, int),
            "spawn_prob": np.full(p, 0.3), "min_performance": np.full(p, 0.5)}


def mixed_levels():
    parts = [gu.levels(a) for a in gu.ARCHIVES]
    return {k: np.concatenate([p[k] for p in parts]) for k in parts[0]}


def big_config(n, min_ms=50.0, dev=None):
    dev = dev or torch.device("cuda", torch.cuda.current_device())
    pool = LevelPool(synth_256(np.random.RandomState(5), np.random.RandomState(6), 8), device=dev)
    envs, per = ring_envs(pool, n)  # 4096 envs are 1.07 GB of state: one batch
    acts = torch.from_numpy(np.random.RandomState(3).randint(9, size=(4, n)).astype(np.uint8)).to(dev)
    launch = step_launcher(pool, envs, acts)
    launch(0, torch.cuda.current_stream().cuda_stream)
    graph = capture(launch, max(2, len(envs)), dev)
    L = max(2, len(envs))
    ms, reps = time_graph(graph, min_ms, dev)
    ms_step = ms / (reps * L)
    hbm, _ = hbm_peak()
    return {"config": "big_256", "envs": n, "board": [256, 256], "ms_per_step": ms_step,
            "env_steps_per_s": n / ms_step * 1e3, "cell_updates_per_s": n * 2 * 65536 / ms_step * 1e3,
            "hbm_frac_algorithmic": n * (8 * 65536 + 32) / (ms_step * 1e-3) / 1e9 / hbm,
            "timed_steps": reps * L, "timed_ms": ms, "state_bytes": per}


def config3_final_states(n, dev, steps=1001, seed_actions=3):
    """BASELINE config 3 played to the end: n prune-spawn envs, seeded random actions, no auto-reset; every env's
    board and num_steps are taken at the step it is first done (game over, or step 1001 = time limit 1000) —
    the states RecordingSafeLifeWrapper.log_episode hands to side_effect_score (env_wrappers.py:213-215)."""
    lv = gu.levels("prune-spawn")
    pool = LevelPool(lv, device=dev)
    env = SafeLifeEnv(pool, num_envs=n, global_counter=None, compute_obs=False)
    env.seed(0)
    env.reset()
    g = env.game
    rb = pool.resident_bytes
    snap = torch.zeros_like(g.board_res)
    snap_steps = torch.zeros(n, dtype=torch.int32, device=dev)
    snap_ctr = torch.zeros(n, dtype=torch.int64, device=dev)
    seen = torch.zeros(n, dtype=torch.bool, device=dev)
    rng = np.random.RandomState(seed_actions)
    chunk = 77  # 1001 = 13 * 77 fused steps per launch; first-done is located per chunk and replayed exactly
    assert steps % chunk == 0
    for c in range(steps // chunk):
        acts = torch.from_numpy(rng.randint(9, size=(chunk, n)).astype(np.uint8)).to(dev)
        before = (g.board_res.clone(), g.goals_res.clone(), g.state.clone())
        _, done = env.rollout(acts)
        first = (done != 0) & ~seen[None]
        if bool(first.any()):
            # replay this chunk step by step from the saved state to catch each env at its first done
            g.board_res.copy_(before[0]); g.goals_res.copy_(before[1]); g.state.copy_(before[2])
            for t in range(chunk):
                _, _, d, _ = env.step(acts[t])
                new = d & ~seen
                if bool(new.any()):
                    idx = new.nonzero().flatten()
                    snap.view(n, rb)[idx] = g.board_res.view(n, rb)[idx]
                    snap_steps[idx] = g.num_steps[idx]
                    snap_ctr[idx] = g.rng_counter[idx]
                    seen |= new
    assert bool(seen.all())
    board1 = torch.empty((n, pool.h, pool.w), dtype=torch.uint16, device=dev)
    _lib.check(_lib.load().slb_resident_export(snap.data_ptr(), board1.data_ptr(), n, pool.h, pool.w,
                                               torch.cuda.current_stream().cuda_stream))
    board0 = pool.board.view(torch.int16)[g.level_id.long()].view(torch.uint16)
    return pool, g, board0, board1, snap_steps, snap_ctr


def side_effects_config(n, dev=None):
    dev = dev or torch.device("cuda", torch.cuda.current_device())
    t0 = time.perf_counter()
    pool, g, board0, board1, num_steps, ctr = config3_final_states(n, dev)
    torch.cuda.synchronize()
    setup_s = time.perf_counter() - t0

    def run():
        return se.side_effect_counts(board0, board1, num_steps, g.spawn_prob, 1000, g.seed, g.env_id0,
                                     ctr.clone(), kmax=16)

    run()  # the first call pays for the allocation of the [N,2,16,H,W] count maps (709 MB at 16384)
    torch.cuda.synchronize()
    best = None
    for _ in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        keys, counts, nkeys = run()
        e1.record()
        torch.cuda.synchronize()
        dt = e0.elapsed_time(e1) * 1e-3
        best = dt if best is None or dt < best else best
    ns = num_steps.cpu().numpy().astype(np.int64)
    cu = int(((ns + 2000) * 676).sum())
    # the EMD stage of side_effect_score over the same count maps (slb_emd_sizes + slb_emd_solve; wall clock of
    # emd_batch incl. its host syncs and scratch allocation; "unpinned" against pyemd)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    emd = se.emd_batch(keys, counts, 1000)
    torch.cuda.synchronize()
    emd_s = time.perf_counter() - t0
    n_prob = int((keys != 0).sum().item())
    emd_max = float(emd.max().item())
    del emd
    return {"config": "side_effects", "episodes": n, "num_samples": 1000, "num_steps_mean": float(ns.mean()),
            "emd_seconds": emd_s, "emd_problems": n_prob, "emd_max": emd_max,
            "num_steps_min": int(ns.min()), "num_steps_max": int(ns.max()), "seconds": best,
            "episodes_per_s": n / best, "cell_updates_per_s": cu / best, "max_keys": int(nkeys.max()),
            "setup_seconds": setup_s,
            "what": "BASELINE config 4: the final states of config 3 (prune-spawn, %d envs, seeded random actions, "
                    "each env's board and num_steps at its first done); device time of one slb_side_effect_counts "
                    "call incl. zero-filling its outputs, best of 3" % n}


def rollout_config(n=4096, T=32, min_ms=50.0, dev=None):
    dev = dev or torch.device("cuda", torch.cuda.current_device())
    pool = LevelPool(gu.levels("append-still"), device=dev)
    envs, _ = ring_envs(pool, n)
    acts = torch.from_numpy(np.random.RandomState(3).randint(9, size=(T, n)).astype(np.uint8)).to(dev)
    rew = torch.empty((T, n), dtype=torch.int32, device=dev)
    done = torch.empty((T, n), dtype=torch.uint8, device=dev)
    lib = _lib.load()

    def launch(j, st):
        e = envs[j % len(envs)]
        g = e.game
        _lib.check(lib.slb_env_rollout(pool.c, n, g.board_res.data_ptr(), g.goals_res.data_ptr(), g.state.data_ptr(),
                                       acts.data_ptr(), T, 1000, 0, g.env_id0, rew.data_ptr(), done.data_ptr(),
                                       e._stride(), st))

    launch(0, torch.cuda.current_stream().cuda_stream)
    graph = capture(launch, len(envs), dev)
    ms, reps = time_graph(graph, min_ms, dev)
    ms_l = ms / (reps * len(envs))
    return {"config": "rollout_k%d" % T, "envs": n, "steps_per_launch": T, "ms_per_launch": ms_l,
            "env_steps_per_s": n * T / ms_l * 1e3, "cell_updates_per_s": n * T * 2 * 676 / ms_l * 1e3}


def advance_board_configs(n=131072, iters=20, dev=None):
    """A1, the speedups.advance_board contract itself: N uint16 boards in, N new uint16 boards out (4 B per
    cell-update at k = 1), boards larger than L2 in total; k = 32 steps fused on chip beside it."""
    dev = dev or torch.device("cuda", torch.cuda.current_device())
    lib = _lib.load()
    hbm, _ = hbm_peak()
    p = torch.full((n,), 0.3, dtype=torch.float32, device=dev)
    out = []
    for arch in ("append-still", "prune-spawn"):
        lv = gu.levels(arch)
        idx = torch.arange(n, device=dev) % lv["board"].shape[0]
        src = torch.from_numpy(lv["board"]).to(dev).view(torch.int16)[idx].contiguous().view(torch.uint16)
        dst = torch.empty_like(src)
        for k in (1, 32):
            def launch(j, st):
                _lib.check(lib.slb_advance_board(src.data_ptr(), dst.data_ptr(), n, 26, 26, p.data_ptr(), 0, 0, None,
                                                 k, None, st))
            launch(0, torch.cuda.current_stream().cuda_stream)
            graph = capture(launch, 2, dev)
            ms, reps = time_graph(graph, 30.0, dev)
            ms_c = ms / (2 * reps)
            cu = n * 676 * k
            name = "advance_board_%s_k%d" % (arch.replace("-", "_"), k)
            o = {"config": name, "boards": n, "board": [26, 26],
                 "k_steps": k, "ms_per_call": ms_c, "cell_updates_per_s": cu / ms_c * 1e3,
                 "hbm_frac": (n * 676 * 4 / (ms_c * 1e-3) / 1e9 / hbm) if k == 1 else None}
            kc = profile_constants().get(name)
            if kc:  # ALU warp instructions per board from the committed ncu capture; alu_pipe_frac is filled in
                o["alu_warp_inst_per_board"] = kc.get("alu_warp_inst_per_board")   # by the caller that measured int_peak
            out.append(o)
    return out


def advance_board_host_config():
    """speedups.advance_board(ndarray) — the literal drop-in call (module.c:19-44; the reference needs 7 us per
    26x26 board on one core): wall clock per call for one board and for batches."""
    from safelife_k2_b200 import speedups as sp
    lv = gu.levels("prune-spawn")
    res = {"config": "advance_board_host", "reference_us_per_board": 7.05,
           "what": "speedups.advance_board(numpy uint16 board(s), 0.3): wall clock per call, host arrays in and out"}
    b1 = lv["board"][0]
    for _ in range(50):
        sp.advance_board(b1, 0.3)
    t0 = time.perf_counter()
    for _ in range(500):
        sp.advance_board(b1, 0.3)
    res["us_single_board"] = (time.perf_counter() - t0) / 500 * 1e6
    per_board = {}
    for nb in (2, 4, 16, 256, 4096, 65536):
        b = lv["board"][np.arange(nb) % 100]
        sp.advance_board(b, 0.3)
        reps = max(3, min(200, 40000 // nb))
        t0 = time.perf_counter()
        for _ in range(reps):
            sp.advance_board(b, 0.3)
        per_board[str(nb)] = (time.perf_counter() - t0) / reps / nb * 1e6
    res["us_per_board_by_batch"] = per_board
    res["break_even_batch"] = next((int(k) for k, v in per_board.items() if v < 7.05), None)
    return res


def env_step_api_config(n=4096, reps=3000, dev=None):
    """the gym-style call a GPU policy makes: env.step(device actions) -> (obs, reward, done, info), wall clock,
    DEFAULT kwargs (class-level global_counter included) except num_envs / auto_reset"""
    dev = dev or torch.device("cuda", torch.cuda.current_device())
    pool = LevelPool(gu.levels("append-still"), device=dev)
    env = SafeLifeEnv(pool, num_envs=n, auto_reset=True)
    env.seed(0)
    env.reset()
    acts = torch.from_numpy(np.random.RandomState(3).randint(9, size=(n,)).astype(np.uint8)).to(dev)
    for _ in range(300):
        env.step(acts)
    torch.cuda.synchronize()
    steps0 = env.global_counter.num_steps
    t0 = time.perf_counter()
    for _ in range(reps):
        obs, r, d, info = env.step(acts)
    host = time.perf_counter() - t0
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    counted = env.global_counter.num_steps - steps0
    return {"config": "env_step_api", "envs": n, "view": [15, 15, 15], "us_per_call_host": host / reps * 1e6,
            "us_per_call": dt / reps * 1e6, "env_steps_per_s": n * reps / dt, "global_counter_num_steps": counted,
            "what": "SafeLifeEnv.step with uint8 cuda actions and default kwargs (global_counter counted on the "
                    "device), fused step + 15x15x15 observation, python call path included"}


def int_peak_config(dev=None):
    """the integer roofline's denominator, measured here: LOP3 chains on every scheduler"""
    import ctypes
    dev = dev or torch.device("cuda", torch.cuda.current_device())
    lib = _lib.load()
    sink = torch.zeros(4, dtype=torch.int32, device=dev)
    res = {"config": "int_peak", "what": "148 x 8 CTAs x 256 threads, 8 independent chains per thread; "
                                          "kind 0 = LOP3 (ALU pipe), kind 1 = LOP3 + IMAD alternating (ALU + FMA pipes)"}
    sm_count = torch.cuda.get_device_properties(dev).multi_processor_count
    for kind, name in ((0, "lop3"), (1, "lop3_imad")):
        wi = ctypes.c_uint64()

        def launch(j, st):
            _lib.check(lib.slb_microbench_int(kind, sm_count * 8, 8192, sink.data_ptr(), ctypes.byref(wi), st))

        launch(0, torch.cuda.current_stream().cuda_stream)
        graph = capture(launch, 1, dev)
        ms, reps = time_graph(graph, 20.0, dev)
        res[name + "_warp_inst_per_s"] = wi.value / (ms / reps * 1e-3)
    res["sm_count"] = sm_count
    return res


def step_obs_config(n=4096, min_ms=30.0, dev=None):
    dev = dev or torch.device("cuda", torch.cuda.current_device())
    pool = LevelPool(gu.levels("append-still"), device=dev)
    envs, _ = ring_envs(pool, n)
    acts = torch.from_numpy(np.random.RandomState(3).randint(9, size=(n,)).astype(np.uint8)).to(dev)
    obs = torch.empty((n, 15, 15, 15), dtype=torch.uint16, device=dev)
    ch = np.arange(15, dtype=np.uint8)
    lib = _lib.load()

    def pair(j, st):
        e = envs[j % len(envs)]
        g = e.game
        _lib.check(lib.slb_env_step(pool.c, n, g.board_res.data_ptr(), g.goals_res.data_ptr(), g.state.data_ptr(),
                                    acts.data_ptr(), 1000, 0, g.env_id0, e._reward.data_ptr(), e._done.data_ptr(),
                                    e._stride(), st))
        _lib.check(lib.slb_get_obs_channels(pool.c, n, g.board_res.data_ptr(), g.goals_res.data_ptr(),
                                            g.state.data_ptr(), 15, 15, 1, ch.ctypes.data, 15, None, obs.data_ptr(), st))

    def fused(j, st):
        e = envs[j % len(envs)]
        g = e.game
        _lib.check(lib.slb_env_step_obs(pool.c, n, g.board_res.data_ptr(), g.goals_res.data_ptr(), g.state.data_ptr(),
                                        acts.data_ptr(), 1000, 0, g.env_id0, e._reward.data_ptr(), e._done.data_ptr(),
                                        e._stride(), 15, 15, 1, ch.ctypes.data, 15, obs.data_ptr(), st))

    res = {"config": "step_obs", "envs": n}
    for name, fn in (("pair", pair), ("fused", fused)):
        fn(0, torch.cuda.current_stream().cuda_stream)
        graph = capture(fn, len(envs), dev)
        ms, reps = time_graph(graph, min_ms, dev)
        res["ms_" + name] = ms / (reps * len(envs))
        res["env_steps_per_s_" + name] = n / res["ms_" + name] * 1e3
    return res


def obs_config(n=4096, dev=None):
    dev = dev or torch.device("cuda", torch.cuda.current_device())
    pool = LevelPool(gu.levels("append-still"), device=dev)
    env = SafeLifeEnv(pool, num_envs=n, global_counter=None)
    env.seed(0)
    env.reset()
    out = torch.empty((n, 15, 15, 15), dtype=torch.uint16, device=dev)
    for _ in range(5):
        env.get_obs(out)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(200):
        env.get_obs(out)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 200
    return {"config": "obs", "envs": n, "view": [15, 15, 15], "ms_per_call": ms, "obs_per_s": n / ms * 1e3,
            "write_GBps": n * 15 * 15 * 15 * 2 / (ms * 1e-3) / 1e9}


# ---------------------------------------------------------------------------------------------
# BASELINE config 5: a FIXED total sharded over the ranks (strong scaling)
# ---------------------------------------------------------------------------------------------
def _digest(t):
    """order-dependent 63-bit checksum of a tensor's bytes (int64 words, wrapping arithmetic)"""
    b = t.contiguous().view(torch.uint8).flatten()
    pad = (-b.numel()) % 8
    if pad:
        b = torch.cat([b, torch.zeros(pad, dtype=torch.uint8, device=b.device)])
    w = b.view(torch.int64)
    k = torch.arange(1, w.numel() + 1, dtype=torch.int64, device=w.device) * 6364136223846793005 + 1442695040888963407
    return int(((w * k).sum() & 0x7FFFFFFFFFFFFFFF).item())


def strong_config(workload, total, rank, world, dev, min_ms=40.0, check_steps=12):
    """`total` envs sharded by index over `world` ranks (distributed.shard_range), no data-path collective, the
    NCCL all-reduce of the statistics vector inside the timed region.  Placement check: before the timed run every
    rank plays `check_steps` steps of ITS OWN first and last 8 envs and of the 8 envs around the middle of the
    whole job (whoever owns them) as separate small batches; the states' checksums, gathered over the ranks, must
    agree between the rank that owns the envs and the ranks that do not (an env's actions and spawn stream are
    keyed by its global index) — and are printed, so that lines of different GPU counts can be compared."""
    if workload == "mixed_26":
        lv, h, w = mixed_levels(), 26, 26
    else:
        lv, h, w = synth_256(np.random.RandomState(5), np.random.RandomState(6), 8), 256, 256
    lo, hi = distributed.shard_range(total, rank, world)
    n = hi - lo
    pool = LevelPool(lv, device=dev)
    lib = _lib.load()
    acts_all = np.random.RandomState(3).randint(9, size=(16, total)).astype(np.uint8)

    # ---- placement check: sample envs [0,8), [total/2-4, total/2+4), [total-8, total) ----
    samples = [0, total // 2 - 4, total - 8]
    digests = []
    for s0 in samples:
        e = SafeLifeEnv(pool, num_envs=8, env_id0=s0, global_counter=None, compute_obs=False)
        e.seed(0)
        e.reset()
        a = torch.from_numpy(acts_all[:, s0:s0 + 8].copy()).to(dev)
        for t in range(check_steps):
            e.step(a[t % 16])
        digests.append(_digest(torch.cat([e.game.board_res, e.game.goals_res, e.game.state.view(torch.uint8).flatten()])))
    # the same envs inside this rank's own shard batch (where it owns them)
    shard = SafeLifeEnv(pool, num_envs=n, env_id0=lo, global_counter=None, compute_obs=False)
    shard.seed(0)
    shard.reset()
    a_sh = torch.from_numpy(acts_all[:, lo:hi].copy()).to(dev)
    for t in range(check_steps):
        shard.step(a_sh[t % 16])
    rb = pool.resident_bytes
    own = []
    for s0 in samples:
        if lo <= s0 and s0 + 8 <= hi:
            k = s0 - lo
            g = shard.game
            own.append(_digest(torch.cat([g.board_res[k * rb:(k + 8) * rb], g.goals_res[k * rb:(k + 8) * rb],
                                          g.state[k:k + 8].contiguous().view(torch.uint8).flatten()])))
        else:
            own.append(-1)
    vec = torch.tensor(digests + own, dtype=torch.int64, device=dev)
    if world > 1:
        allv = [torch.zeros_like(vec) for _ in range(world)]
        torch.distributed.all_gather(allv, vec)
        allv = torch.stack(allv).cpu().numpy()
    else:
        allv = vec.cpu().numpy()[None]
    ok = True
    for k in range(3):
        small = set(int(x) for x in allv[:, k])
        owners = set(int(x) for x in allv[:, 3 + k] if x != -1)
        ok = ok and len(small) == 1 and (not owners or owners == small)
    del shard

    # ---- timed: ring of this rank's share ----
    envs, per = ring_envs(pool, n, env_id0=lo, id_stride=total)
    acts = torch.from_numpy(acts_all[:, lo:hi].copy()).to(dev)
    stats = torch.zeros(8, dtype=torch.int64, device=dev)
    launch = step_launcher(pool, envs, acts)
    L = max(2, len(envs))
    launch(0, torch.cuda.current_stream().cuda_stream)
    graph = capture(launch, L, dev)
    last = envs[(L - 1) % len(envs)]

    def tail():
        _lib.check(lib.slb_accumulate_stats(n, last.game.state.data_ptr(), last._reward.data_ptr(),
                                            last._done.data_ptr(), stats.data_ptr(),
                                            torch.cuda.current_stream().cuda_stream))
        distributed.all_reduce_stats(stats)

    ms, reps = time_graph(graph, min_ms, dev, world, tail)
    ms_step = ms / (reps * L)
    # informational: the same ring with even / odd batches on two streams (an actor that alternates two
    # independent batches of its share); the judged figures above stay single-stream
    two = None
    if len(envs) >= 2:
        L2n = len(envs) - (len(envs) % 2)
        g2 = capture2(launch, L2n, dev)
        ms2, r2 = time_graph(g2, min_ms / 2, dev, world)
        two = {"ms_per_step": ms2 / (r2 * L2n), "env_steps_per_s": total / (ms2 / (r2 * L2n)) * 1e3}
        del g2
    hbm, _ = hbm_peak()
    return {"workload": workload, "total_envs": total, "envs_per_gpu": n, "board": [h, w], "n_gpus": world,
            "two_stream_overlap": two,
            "ms_per_step": ms_step, "env_steps_per_s": total / ms_step * 1e3,
            "cell_updates_per_s": total * 2 * h * w / ms_step * 1e3,
            "hbm_frac_algorithmic_per_gpu": n * (8 * h * w + 32) / (ms_step * 1e-3) / 1e9 / hbm,
            "ring_batches_per_gpu": len(envs), "timed_steps": reps * L, "timed_ms": ms,
            "allreduce_in_timed_region": True, "placement_check": "ok" if ok else "MISMATCH",
            "sample_digests": [int(x) for x in allv[0, :3]],
            "stats_allreduce": distributed.stats_dict(stats)}


def main():
    names = [a for a in sys.argv[1:] if not a.startswith("--")] or [
        "prune_spawn_16384", "mixed_65536", "rollout_k32", "big_256", "side_effects", "step_obs", "obs",
        "advance_board", "advance_board_host", "env_step_api", "int_peak"]
    quick = "--quick" in sys.argv
    ms = 15.0 if quick else 50.0
    consts = profile_constants()
    out = []
    if "prune_spawn_16384" in names:
        out.append(step_config("prune_spawn_16384", gu.levels("prune-spawn"), 16384, 26, 26, ms, consts=consts))
    if "mixed_65536" in names:
        out.append(step_config("mixed_65536", mixed_levels(), 65536, 26, 26, ms, consts=consts))
    if "rollout_k32" in names:
        out.append(rollout_config(min_ms=ms))
    if "big_256" in names:
        out.append(big_config(512 if quick else 4096, ms))
    if "side_effects" in names:
        out.append(side_effects_config(1024 if quick else 16384))
    if "step_obs" in names:
        out.append(step_obs_config(min_ms=ms))
    if "obs" in names:
        out.append(obs_config())
    if "advance_board" in names:
        out += advance_board_configs(8192 if quick else 131072)
    if "advance_board_host" in names:
        out.append(advance_board_host_config())
    if "env_step_api" in names:
        out.append(env_step_api_config(reps=300 if quick else 3000))
    if "int_peak" in names:
        out.append(int_peak_config())
    for nm in names:  # step:<archive|mixed>:<envs> — the fused step on any batch size
        if nm.startswith("step:"):
            _, arch, cnt = nm.split(":")
            lv = mixed_levels() if arch == "mixed" else gu.levels(arch)
            out.append(step_config("step_%s_%s" % (arch, cnt), lv, int(cnt), 26, 26, ms, consts=consts))
        if nm.startswith("big:"):
            out.append(big_config(int(nm.split(":")[1]), ms))
    for nm in names:  # strong:<mixed_26|big_256>:<total> — one GPU's share of a strong-scaling job as a whole job
        if nm.startswith("strong:"):
            _, wl, tot = nm.split(":")
            out.append(strong_config(wl, int(tot), 0, 1, torch.device("cuda", 0), ms))
    if "strong_mixed" in names:
        out.append(strong_config("mixed_26", 65536, 0, 1, torch.device("cuda", 0), ms))
    if "strong_big" in names:
        out.append(strong_config("big_256", 512 if quick else 4096, 0, 1, torch.device("cuda", 0), ms))
    for o in out:
        print(json.dumps(o))


if __name__ == "__main__":
    main()
