"""Loaders for the committed golden fixtures (tests/golden/*.npz, made by make_golden.py)."""
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
ARCHIVES = ["append-still", "prune-still", "append-dynamic", "prune-dynamic", "append-spawn",
            "prune-spawn", "navigation", "prune-still-hard"]


def load(name):
    with np.load(os.path.join(GOLDEN, name)) as d:
        return {k: d[k] for k in d.files}


_levels = None


def levels(archive):
    """dict of stacked arrays (board, goals, agent_loc, orientation, spawn_prob, min_performance)"""
    global _levels
    if _levels is None:
        _levels = load("levels_v1.npz")
    key = archive.replace("-", "_")
    return {f: _levels["%s__%s" % (key, f)] for f in
            ("board", "goals", "agent_loc", "orientation", "spawn_prob", "min_performance")}


def level_slice(lv, lo, hi):
    return {k: v[lo:hi] for k, v in lv.items()}


def kat_cases():
    d = load("kat_advance.npz")
    seed = int(d["seed"])
    for env, name in enumerate(d["names"]):
        yield str(name), env, seed, d["in_%d" % env], d["out_%d" % env], float(d["p_%d" % env])
