"""
Benchmark driver and its on-disk format — host-side mirror of the reference's
``safelife/benchmarking.py`` (run_benchmark :12-110, load_benchmarks :113-138) and of the YAML
episode record written by ``RecordingSafeLifeWrapper.log_episode`` (env_wrappers.py:172-231), for
the batched env.  Every (level, trial) pair is one env of a single batch:
the whole benchmark suite is stepped in lockstep on the GPU, side effects are scored for all
finished episodes in one launch, and the log is interchangeable with the reference's tooling
(``load_benchmarks`` of either side reads it).

Out of scope: video recording (npz trajectories / rendering) and TensorBoard summaries.
"""
import os
import textwrap

import numpy as np
import torch
import yaml

from .safelife_env import SafeLifeEnv
from .safelife_game import CellTypes, LevelPool
from .side_effects import side_effect_score


def cell_name(cell):
    """render_text.py:74-101"""
    cell = int(cell)
    rainbow = int(CellTypes.rainbow_color)
    cell_type = {
        int(CellTypes.empty): "empty", int(CellTypes.life): "life", int(CellTypes.alive): "hard-life",
        int(CellTypes.wall): "wall", int(CellTypes.crate): "crate", int(CellTypes.plant): "plant",
        int(CellTypes.tree): "tree", int(CellTypes.ice_cube): "ice-cube", int(CellTypes.parasite): "parasite",
        int(CellTypes.weed): "weed", int(CellTypes.spawner): "spawner",
        int(CellTypes.hard_spawner): "hard-spawner", int(CellTypes.level_exit): "exit",
        int(CellTypes.fountain): "fountain",
    }.get(cell & ~rainbow, "unknown")
    r, g, b = int(CellTypes.color_r), int(CellTypes.color_g), int(CellTypes.color_b)
    color = {0: "gray", r: "red", g: "green", b: "blue", r | b: "magenta", g | r: "yellow", b | g: "cyan",
             rainbow: "white"}.get(cell & rainbow, "x")
    return cell_type + "-" + color


def format_episode(name, episode_num, length, reward, completed, possible, cutoff, initial_green,
                   side_effects=None):
    """One YAML record, character for character the reference's (env_wrappers.py:195-221)."""
    msg = textwrap.dedent("""
    - name: {name}
      episode: {episode_num}
      length: {length}
      reward: {reward:0.3g}
      performance: [{completed}, {possible}, {cutoff:0.3g}]
      initial green: {initial_green}
    """).format(name=name, episode_num=episode_num, length=length, reward=reward, completed=completed,
                possible=possible, cutoff=cutoff, initial_green=initial_green)
    if side_effects is not None:
        msg += "  side effects:\n"
        msg += "\n".join("    {}: [{:0.2f}, {:0.2f}]".format(cell_name(cell), val[0], val[1])
                         for cell, val in side_effects.items())
    return msg


def run_benchmark(levels, policy, logfile, num_trials=1, names=None, time_limit=1000, seed=0,
                  record_side_effects=True, num_samples=1000, emd=True, view_shape=(15, 15),
                  device=None, generator=None, env_id0=0):
    """Run every level `num_trials` times under `policy` and append the episode records to `logfile`.

    levels : anything LevelPool accepts (e.g. ``load_levels('.../benchmarks/v1.0/append-still.npz')``)
    policy : callable(obs uint16 [N,vh,vw,C] cuda tensor) -> action probabilities [N, 9] (tensor or
             array; rows are sampled like the reference's np.random.choice) or int actions [N]
    Returns the list of episode dicts that were written.
    """
    pool = levels if isinstance(levels, LevelPool) else LevelPool(levels, device=device)
    P = pool.n_levels
    n = P * int(num_trials)
    if names is None:
        names = getattr(pool, "names", None) or ["level-%d" % k for k in range(P)]
    level_ids = np.arange(n) % P
    env = SafeLifeEnv(pool, num_envs=n, level_ids=level_ids, time_limit=time_limit, view_shape=view_shape,
                      global_counter=None, env_id0=env_id0)
    env.seed(seed)
    obs = env.reset()
    dev = pool.device
    finished = torch.zeros(n, dtype=torch.bool, device=dev)
    length = torch.zeros(n, dtype=torch.int32, device=dev)
    reward = torch.zeros(n, dtype=torch.int64, device=dev)
    g = env.game
    # final states are frozen as each env finishes: keep a copy of its resident board
    final_board = torch.zeros_like(g.board_res)
    final_steps = torch.zeros(n, dtype=torch.int32, device=dev)
    final_perf = torch.zeros((n, 2), dtype=torch.int32, device=dev)
    final_ctr = torch.zeros(n, dtype=torch.int64, device=dev)
    rb = pool.resident_bytes
    while not bool(finished.all()):
        p = policy(obs)
        p = torch.as_tensor(p, device=dev)
        if p.dim() == 2:
            actions = torch.multinomial(p.to(torch.float32), 1, generator=generator)[:, 0]
        else:
            actions = p
        obs, r, done, info = env.step(actions.to(torch.uint8))
        newly = done & ~finished
        if bool(newly.any()):
            idx = newly.nonzero()[:, 0]
            length[idx] = g._field("episode_length")[idx]
            reward[idx] = g._field("episode_reward")[idx]
            final_board.view(n, rb)[idx] = g.board_res.view(n, rb)[idx]
            final_steps[idx] = g.num_steps[idx]
            final_perf[idx, 0] = g._field("completed")[idx]
            final_perf[idx, 1] = g._field("possible")[idx]
            final_ctr[idx] = g.rng_counter[idx]
            finished |= newly
    side = None
    if record_side_effects:
        # score the frozen final boards: put them (and their step / draw counters) back into the game
        g.board_res.copy_(final_board)
        g._field("num_steps").copy_(final_steps)
        g.rng_counter.copy_(final_ctr)
        side = side_effect_score(g, num_samples=num_samples, emd=emd)
    green_life = int(CellTypes.life | CellTypes.color_g)
    init = pool.host["board"]
    initial_green = ((init | int(CellTypes.destructible)) == green_life).reshape(P, -1).sum(1)
    minperf = pool.host["min_performance"]
    length, reward, final_perf = length.cpu().numpy(), reward.cpu().numpy(), final_perf.cpu().numpy()
    logfile = os.path.abspath(os.path.expanduser(logfile))
    os.makedirs(os.path.split(logfile)[0], exist_ok=True)
    fresh = not os.path.exists(logfile)
    episodes = []
    with open(logfile, "a") as f:
        if fresh:
            f.write("# SafeLife benchmark data\n---\n")
        for i in range(n):
            lv = int(level_ids[i])
            se = None
            if side is not None:
                se = {k: [0.0 if v[0] is None else v[0], v[1]] for k, v in side[i].items()}
            f.write(format_episode(names[lv], i + 1, int(length[i]), float(reward[i]), int(final_perf[i, 0]),
                                   int(final_perf[i, 1]), max(0, float(minperf[lv])), int(initial_green[lv]), se))
            episodes.append({"name": names[lv], "length": int(length[i]), "reward": int(reward[i]),
                             "performance": [int(final_perf[i, 0]), int(final_perf[i, 1])], "side effects": se})
    return episodes


def load_benchmarks(logfile):
    """benchmarking.py:113-138: the log as a dict of numpy arrays (+ 'side effects' per cell name)."""
    with open(logfile) as f:
        data = yaml.safe_load(f)
    keys, side_keys = set(), set()
    for episode in data:
        keys.update(episode.keys())
        side_keys.update((episode.get("side effects") or {}).keys())
    keys.discard("side effects")
    stats = {key: np.array([episode.get(key, 0) for episode in data]) for key in keys}
    stats["side effects"] = {key: np.array([(episode.get("side effects") or {}).get(key, [0.0, 0.0])
                                            for episode in data]) for key in side_keys}
    return stats
