#!/usr/bin/env python
"""
bench_configs.py — throughput of the OTHER BASELINE.json configurations on one B200 (bench.py is the
judged line for config 2).  Prints one JSON line per configuration:

  prune_spawn_16384   config 3: prune-spawn 26x26, 16384 envs, stochastic spawners, full env step, k=1
  mixed_65536         config 5a: 65536 envs cycled over all 8 v1.0 archives, k=1
  rollout_k32         config 2 with 32 env steps fused per launch (pre-staged actions)
  big_256             config 5b: 4096 envs at 256x256 (synthetic levels of SURVEY §8(d)), k=1
  side_effects        config 4: side_effect_score counts (1000 samples) for finished prune-spawn episodes
  obs                 get_obs 15x15x15 for 4096 envs
  advance_board       the speedups.advance_board contract: uint16 boards in / out, k = 1 and k = 32
  env_step_api        wall clock of SafeLifeEnv.step (device actions in, observation out) per call
  int_peak            LOP3 / LOP3+IMAD issue-rate microbenchmark (denominator of the integer roofline)

    python bench_configs.py [names...] [--quick]
"""
import json
import sys
import time

import numpy as np
import torch

sys.path.insert(0, __file__.rsplit("/", 1)[0] or ".")
from safelife_k2_b200 import _lib  # noqa: E402
from safelife_k2_b200.safelife_env import SafeLifeEnv  # noqa: E402
from safelife_k2_b200.safelife_game import LevelPool  # noqa: E402
from safelife_k2_b200 import side_effects as se  # noqa: E402
from tests import golden_util as gu  # noqa: E402

HBM = 6458.1
try:
    HBM = float(json.load(open(__file__.rsplit("/", 1)[0] + "/MEASURED_PEAKS.json"))["hbm_gbs"])
except Exception:
    pass


def timed(fn, iters, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters  # ms


def ring_envs(pool, n, min_bytes=2.2 * 126e6, **kw):
    per = n * (2 * pool.resident_bytes + 64)
    ring = max(2, int(np.ceil(min_bytes / per)))
    envs = []
    for r in range(ring):
        e = SafeLifeEnv(pool, num_envs=n, env_id0=r * n, global_counter=None, compute_obs=False, auto_reset=True, **kw)
        e.seed(0)
        e.reset()
        envs.append(e)
    return envs


def step_config(name, lv, n, iters, h, w):
    pool = LevelPool(lv)
    envs = ring_envs(pool, n)
    acts = torch.from_numpy(np.random.RandomState(3).randint(9, size=(16, n)).astype(np.uint8)).cuda()
    lib, st = _lib.load(), torch.cuda.current_stream().cuda_stream
    k = [0]

    def one():
        e = envs[k[0] % len(envs)]
        g = e.game
        _lib.check(lib.slb_env_step(pool.c, n, g.board_res.data_ptr(), g.goals_res.data_ptr(), g.state.data_ptr(),
                                    acts[k[0] % 16].data_ptr(), 1000, 0, g.env_id0, e._reward.data_ptr(),
                                    e._done.data_ptr(), e._stride(), st))
        k[0] += 1

    ms = timed(one, iters)
    bytes_alg = n * (8 * h * w + 32)
    return {"config": name, "envs": n, "board": [h, w], "ms_per_step": ms, "env_steps_per_s": n / ms * 1e3,
            "cell_updates_per_s": n * 2 * h * w / ms * 1e3, "hbm_frac_algorithmic": bytes_alg / (ms * 1e-3) / 1e9 / HBM,
            "ring": len(envs)}


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


def main():
    names = [a for a in sys.argv[1:] if not a.startswith("--")] or ["prune_spawn_16384", "mixed_65536", "rollout_k32",
                                                                   "big_256", "side_effects", "step_obs", "obs", "advance_board", "env_step_api", "int_peak"]
    quick = "--quick" in sys.argv
    out = []
    if "prune_spawn_16384" in names:
        out.append(step_config("prune_spawn_16384", gu.levels("prune-spawn"), 16384, 100 if quick else 400, 26, 26))
    if "mixed_65536" in names:
        parts = [gu.levels(a) for a in gu.ARCHIVES]
        lv = {k: np.concatenate([p[k] for p in parts]) for k in parts[0]}
        out.append(step_config("mixed_65536", lv, 65536, 50 if quick else 200, 26, 26))
    if "rollout_k32" in names:
        pool = LevelPool(gu.levels("append-still"))
        n, T = 4096, 32
        envs = ring_envs(pool, n)
        acts = torch.from_numpy(np.random.RandomState(3).randint(9, size=(T, n)).astype(np.uint8)).cuda()
        rew = torch.empty((T, n), dtype=torch.int32, device="cuda")
        done = torch.empty((T, n), dtype=torch.uint8, device="cuda")
        k = [0]

        def roll():
            envs[k[0] % len(envs)].rollout(acts, rew, done)
            k[0] += 1

        ms = timed(roll, 50 if quick else 200)
        out.append({"config": "rollout_k32", "envs": n, "steps_per_launch": T, "ms_per_launch": ms,
                    "env_steps_per_s": n * T / ms * 1e3, "cell_updates_per_s": n * T * 2 * 676 / ms * 1e3})
    if "big_256" in names:
        n = 512 if quick else 4096
        lv = synth_256(np.random.RandomState(5), np.random.RandomState(6), 8)
        pool = LevelPool(lv)
        env = SafeLifeEnv(pool, num_envs=n, global_counter=None, compute_obs=False, auto_reset=True)
        env.seed(0)
        env.reset()
        acts = torch.from_numpy(np.random.RandomState(3).randint(9, size=(n,)).astype(np.uint8)).cuda()
        lib, st = _lib.load(), torch.cuda.current_stream().cuda_stream
        g = env.game

        def one():
            _lib.check(lib.slb_env_step(pool.c, n, g.board_res.data_ptr(), g.goals_res.data_ptr(), g.state.data_ptr(),
                                        acts.data_ptr(), 1000, 0, 0, env._reward.data_ptr(), env._done.data_ptr(),
                                        env._stride(), st))

        ms = timed(one, 5 if quick else 20, warm=2)
        out.append({"config": "big_256", "envs": n, "board": [256, 256], "ms_per_step": ms,
                    "env_steps_per_s": n / ms * 1e3, "cell_updates_per_s": n * 2 * 65536 / ms * 1e3,
                    "hbm_frac_algorithmic": n * (8 * 65536 + 32) / (ms * 1e-3) / 1e9 / HBM})
    if "side_effects" in names:
        n = 1024 if quick else 16384
        lv = gu.levels("prune-spawn")
        pool = LevelPool(lv)
        env = SafeLifeEnv(pool, num_envs=n, global_counter=None, compute_obs=False)
        env.seed(0)
        env.reset()
        T = 100
        acts = torch.from_numpy(np.random.RandomState(3).randint(9, size=(T, n)).astype(np.uint8)).cuda()
        env.rollout(acts)
        g = env.game
        lvl = g.level_id.long()
        board0 = pool.board.view(torch.int16)[lvl].view(torch.uint16)
        board1 = g.board
        def run():
            return se.side_effect_counts(board0, board1, g.num_steps.contiguous(), g.spawn_prob, 1000,
                                         g.seed, g.env_id0, None, kmax=16)

        run()  # the first call pays for the allocation of the [N,2,16,H,W] count maps (709 MB)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        keys, counts, nkeys = run()
        e1.record()
        torch.cuda.synchronize()
        wall = time.perf_counter() - t0
        dt = e0.elapsed_time(e1) * 1e-3
        cu = n * (T + 2000) * 676
        out.append({"config": "side_effects", "episodes": n, "num_steps": T, "num_samples": 1000, "seconds": dt,
                    "wall_seconds": wall, "episodes_per_s": n / dt, "cell_updates_per_s": cu / dt,
                    "max_keys": int(nkeys.max()),
                    "what": "prune-spawn episodes (spawners: both boards change at every sample); device time of "
                            "one slb_side_effect_counts call incl. zero-filling the outputs"})
        # the same for episodes of a still archive: the inaction board is a fixed point from the start
        lv2 = gu.levels("append-still")
        pool2 = LevelPool(lv2)
        env2 = SafeLifeEnv(pool2, num_envs=n, global_counter=None, compute_obs=False)
        env2.seed(0)
        env2.reset()
        env2.rollout(acts)
        g2 = env2.game
        b0 = pool2.board.view(torch.int16)[g2.level_id.long()].view(torch.uint16)
        b1 = g2.board

        def run2():
            return se.side_effect_counts(b0, b1, g2.num_steps.contiguous(), g2.spawn_prob, 1000, g2.seed,
                                         g2.env_id0, None, kmax=16)

        run2()
        torch.cuda.synchronize()
        e0.record()
        run2()
        e1.record()
        torch.cuda.synchronize()
        dt2 = e0.elapsed_time(e1) * 1e-3
        out.append({"config": "side_effects_still", "episodes": n, "num_steps": T, "num_samples": 1000,
                    "seconds": dt2, "episodes_per_s": n / dt2, "nominal_cell_updates_per_s": cu / dt2,
                    "what": "append-still episodes after 100 random actions: runs of identical states are counted "
                            "once, draw-free fixed points end the rollout"})
    if "step_obs" in names:
        # the policy-in-the-loop pair: fused step, then the observation for the next action, one ring pass as a
        # CUDA graph (device time; the Python call path costs more than both kernels at this size)
        pool = LevelPool(gu.levels("append-still"))
        n = 4096
        envs = ring_envs(pool, n)
        acts = torch.from_numpy(np.random.RandomState(3).randint(9, size=(n,)).astype(np.uint8)).cuda()
        obs = torch.empty((n, 15, 15, 15), dtype=torch.uint16, device="cuda")
        ch = np.arange(15, dtype=np.uint8)
        lib = _lib.load()
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.stream(side):
            st = side.cuda_stream
            with torch.cuda.graph(graph, stream=side):
                for e in envs:
                    g = e.game
                    _lib.check(lib.slb_env_step(pool.c, n, g.board_res.data_ptr(), g.goals_res.data_ptr(),
                                                g.state.data_ptr(), acts.data_ptr(), 1000, 0, g.env_id0,
                                                e._reward.data_ptr(), e._done.data_ptr(), e._stride(), st))
                    _lib.check(lib.slb_get_obs_channels(pool.c, n, g.board_res.data_ptr(), g.goals_res.data_ptr(),
                                                        g.state.data_ptr(), 15, 15, 1, ch.ctypes.data, 15, None,
                                                        obs.data_ptr(), st))
        graph2 = torch.cuda.CUDAGraph()
        with torch.cuda.stream(side):
            st = side.cuda_stream
            with torch.cuda.graph(graph2, stream=side):
                for e in envs:
                    g = e.game
                    _lib.check(lib.slb_env_step_obs(pool.c, n, g.board_res.data_ptr(), g.goals_res.data_ptr(),
                                                    g.state.data_ptr(), acts.data_ptr(), 1000, 0, g.env_id0,
                                                    e._reward.data_ptr(), e._done.data_ptr(), e._stride(), 15, 15, 1,
                                                    ch.ctypes.data, 15, obs.data_ptr(), st))
        torch.cuda.synchronize()
        ms = timed(graph.replay, 20 if quick else 100) / len(envs)
        ms2 = timed(graph2.replay, 20 if quick else 100) / len(envs)
        out.append({"config": "step_obs", "envs": n, "ms_per_step_plus_obs": ms, "env_steps_per_s": n / ms * 1e3,
                    "ms_fused_step_obs": ms2, "env_steps_per_s_fused": n / ms2 * 1e3,
                    "what": "slb_env_step + slb_get_obs_channels (15x15x15) per env batch vs slb_env_step_obs, ring of %d batches" % len(envs)})
    if "obs" in names:
        pool = LevelPool(gu.levels("append-still"))
        n = 4096
        env = SafeLifeEnv(pool, num_envs=n, global_counter=None)
        env.seed(0)
        env.reset()
        ms = timed(env.get_obs, 200)
        out.append({"config": "obs", "envs": n, "view": [15, 15, 15], "ms_per_call": ms, "obs_per_s": n / ms * 1e3,
                    "write_GBps": n * 15 * 15 * 15 * 2 / (ms * 1e-3) / 1e9})
    if "advance_board" in names:
        # A1, the speedups.advance_board contract itself: N uint16 boards in, N new uint16 boards out
        # (4 B per cell-update at k = 1), boards larger than L2 in total; k = 32 steps fused on chip beside it
        n = 8192 if quick else 131072
        lib, st = _lib.load(), torch.cuda.current_stream().cuda_stream
        p = torch.full((n,), 0.3, dtype=torch.float32, device="cuda")
        for arch in ("append-still", "prune-spawn"):
            lv = gu.levels(arch)
            idx = torch.arange(n, device="cuda") % lv["board"].shape[0]
            src = torch.from_numpy(lv["board"]).cuda().view(torch.int16)[idx].contiguous().view(torch.uint16)
            dst = torch.empty_like(src)
            for k in (1, 32):
                def go():
                    _lib.check(lib.slb_advance_board(src.data_ptr(), dst.data_ptr(), n, 26, 26, p.data_ptr(), 0, 0, None,
                                                     k, None, st))
                ms = timed(go, 5 if quick else 20, warm=3)
                cu = n * 676 * k
                out.append({"config": "advance_board_%s_k%d" % (arch, k), "boards": n, "board": [26, 26], "k_steps": k,
                            "ms_per_call": ms, "cell_updates_per_s": cu / ms * 1e3,
                            "hbm_frac": (n * 676 * 4 / (ms * 1e-3) / 1e9 / HBM) if k == 1 else None,
                            "what": "slb_advance_board (the speedups.advance_board contract: uint16 boards in, new uint16 "
                                    "boards out), one stream per board"})
    if "env_step_api" in names:
        # the gym-style call a GPU policy makes: env.step(device actions) -> (obs, reward, done, info), wall clock
        pool = LevelPool(gu.levels("append-still"))
        n = 4096
        env = SafeLifeEnv(pool, num_envs=n, global_counter=None, auto_reset=True)
        env.seed(0)
        env.reset()
        acts = torch.from_numpy(np.random.RandomState(3).randint(9, size=(n,)).astype(np.uint8)).cuda()
        for _ in range(200):
            env.step(acts)
        torch.cuda.synchronize()
        reps = 300 if quick else 3000
        t0 = time.perf_counter()
        for _ in range(reps):
            obs, r, d, info = env.step(acts)
        host = time.perf_counter() - t0
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        out.append({"config": "env_step_api", "envs": n, "view": [15, 15, 15], "us_per_call_host": host / reps * 1e6,
                    "us_per_call": dt / reps * 1e6, "env_steps_per_s": n * reps / dt,
                    "what": "SafeLifeEnv.step with uint8 cuda actions, fused step + observation, python call path included"})
    if "int_peak" in names:
        # the integer roofline's denominator, measured here: LOP3 chains on every scheduler
        import ctypes
        lib = _lib.load()
        sink = torch.zeros(4, dtype=torch.int32, device="cuda")
        res = {"config": "int_peak", "what": "148 x 8 CTAs x 256 threads, 8 independent chains per thread; "
                                              "kind 0 = LOP3 (ALU pipe), kind 1 = LOP3 + IMAD alternating (ALU + FMA pipes)"}
        sm_count = torch.cuda.get_device_properties(0).multi_processor_count
        for kind, name in ((0, "lop3"), (1, "lop3_imad")):
            wi = ctypes.c_uint64()
            st = torch.cuda.current_stream().cuda_stream

            def go():
                _lib.check(lib.slb_microbench_int(kind, sm_count * 8, 8192, sink.data_ptr(), ctypes.byref(wi), st))

            ms = timed(go, 10, warm=3)
            res[name + "_warp_inst_per_s"] = wi.value / (ms * 1e-3)
            res[name + "_ms"] = ms
        try:
            import subprocess
            mhz = float(subprocess.run(["nvidia-smi", "-i", "0", "--query-gpu=clocks.sm", "--format=csv,noheader,nounits"],
                                       capture_output=True, text=True).stdout.strip())
            res["sm_mhz_after"] = mhz
        except Exception:
            pass
        res["sm_count"] = sm_count
        out.append(res)
    for o in out:
        print(json.dumps(o))


if __name__ == "__main__":
    main()
