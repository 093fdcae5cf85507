"""How much of the multi-wave launch time of the mixed archives is tail from the variance of the cost per env?
One 8192-env batch whose env i plays level i of an 8192-entry pool: (a) levels in random order, (b) sorted so
that the expensive archives come first (longest processing time first), (c) cheapest first."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import bench_configs as bc
from tests import golden_util as gu

n = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
cost = {}
for a in gu.ARCHIVES:
    r = bc.step_config("step_%s" % a, gu.levels(a), n, 26, 26, 20.0)
    cost[a] = r["ms_per_step"]
    print(json.dumps({"archive": a, "ms_per_step": r["ms_per_step"]}), flush=True)
parts = [gu.levels(a) for a in gu.ARCHIVES]
per = n // len(parts)
rng = np.random.RandomState(0)
order = sorted(range(len(parts)), key=lambda k: -cost[gu.ARCHIVES[k]])


def pool_of(seq):
    out = {}
    for key in parts[0]:
        out[key] = np.concatenate([parts[k][key][np.arange(per) % parts[k][key].shape[0]] for k in seq])
    return out


def shuffled(lv):
    p = rng.permutation(n)
    return {k: v[p] for k, v in lv.items()}


for name, lv in (("expensive_first", pool_of(order)), ("cheap_first", pool_of(order[::-1])),
                 ("shuffled", shuffled(pool_of(order)))):
    r = bc.step_config(name, lv, n, 26, 26, 30.0)
    print(json.dumps({"order": name, "ms_per_step": r["ms_per_step"], "env_steps_per_s": r["env_steps_per_s"]}), flush=True)
