"""
Side-effect score — host-side mirror of the reference's ``safelife/side_effects.py``.

The expensive part (side_effects.py:131-141: catch-up of the inaction board, 2 x num_samples
paired advance_board steps, per-cell-type count maps of _add_cell_distribution :59-86) runs on the
GPU for all finished episodes at once (slb_side_effect_counts).  The host then does what the
reference does with the maps: divide by n (:89-92), np.sum, and the earth-mover distance.

EMD: the reference calls third-party ``pyemd.emd`` (pyemd 0.5.1, not vendored, not installed
here) at side_effects.py:56.  If pyemd is importable it is used; otherwise the same transportation
problem (ground distance tanh(manhattan_torus/5), extra_mass_penalty 1.0, :44-56) is solved with
scipy's LP.  That one number is "parity unpinned" (DESIGN.md); everything before it is bit-exact.
"""
import numpy as np
import torch

from . import _lib


def side_effect_counts(board0, board1, num_steps, spawn_prob, num_samples=1000, seed=0, env_id0=0,
                       counters=None, kmax=32):
    """board0/board1: uint16 [N,H,W] cuda tensors (initial and final boards of N episodes);
    num_steps: int32 [N]; spawn_prob: float32 [N].
    -> keys uint16 [N,kmax] (0 = empty), counts uint16 [N,2,kmax,H,W], nkeys int32 [N]."""
    n, h, w = board0.shape
    dev = board0.device
    keys = torch.zeros((n, kmax), dtype=torch.uint16, device=dev)
    counts = torch.zeros((n, 2, kmax, h, w), dtype=torch.uint16, device=dev)
    nkeys = torch.zeros(n, dtype=torch.int32, device=dev)
    num_steps = torch.as_tensor(num_steps, dtype=torch.int32, device=dev).contiguous()
    spawn_prob = torch.as_tensor(spawn_prob, dtype=torch.float32, device=dev).expand(n).contiguous()
    scratch = None
    if 4 * h * w * 2 > 200 * 1024:
        scratch = torch.empty((n, 4, h, w), dtype=torch.uint16, device=dev)
    with torch.cuda.device(dev):
        _lib.check(_lib.load().slb_side_effect_counts(
            board0.contiguous().data_ptr(), board1.contiguous().data_ptr(), n, h, w,
            num_steps.data_ptr(), int(num_samples), spawn_prob.data_ptr(), int(seed), int(env_id0),
            _lib.ptr(counters), kmax, keys.data_ptr(), counts.data_ptr(), nkeys.data_ptr(),
            _lib.ptr(scratch), _lib.current_stream()))
    if int((nkeys < 0).sum().item()):
        raise _lib.SlbError(_lib.SLB_ERR_OVERFLOW, "more than %d distinct cell types" % kmax)
    return keys, counts, nkeys


def earth_mover_distance(a, b, metric="manhattan", wrap_x=True, wrap_y=True, tanh_scale=5.0,
                         extra_mass_penalty=1.0):
    """side_effects.py:12-56.  UNPINNED when pyemd is absent (see module docstring)."""
    a = np.asanyarray(a, dtype=float)
    b = np.asanyarray(b, dtype=float)
    x, y = np.meshgrid(np.arange(a.shape[1]), np.arange(a.shape[0]))
    delta = np.abs(a - b)
    changed = delta > 1e-3 * np.max(delta)
    if not changed.any():
        return 0.0
    dx = np.subtract.outer(x[changed], x[changed])
    dy = np.subtract.outer(y[changed], y[changed])
    if wrap_x:
        dx = np.minimum(dx, a.shape[1] - dx)
    if wrap_y:
        dy = np.minimum(dy, a.shape[0] - dy)
    if metric == "manhattan":
        dist = (np.abs(dx) + np.abs(dy)).astype(float)
    else:
        dist = np.sqrt(dx * dx + dy * dy)
    if tanh_scale > 0:
        dist = np.tanh(dist / tanh_scale)
    try:
        import pyemd
        return pyemd.emd(a[changed], b[changed], dist, extra_mass_penalty)
    except ImportError:
        return _emd_lp(a[changed], b[changed], dist, extra_mass_penalty)


def _emd_lp(p, q, dist, extra_mass_penalty):
    """min-cost flow of min(sum p, sum q) mass + penalty * |sum p - sum q| (pyemd's definition for
    unequal masses, un-normalised)."""
    from scipy.optimize import linprog

    from scipy.sparse import coo_matrix

    n = len(p)
    if extra_mass_penalty < 0:
        extra_mass_penalty = dist.max()
    # flow f[i, j] >= 0 (variable i * n + j): row sums <= p, column sums <= q, total = min(sum p, sum q);
    # the constraint matrix has 2 n^2 ones — kept sparse (a dense one is 2 n^3 doubles)
    var = np.arange(n * n)
    rows = np.concatenate([var // n, n + var % n])
    A_ub = coo_matrix((np.ones(2 * n * n), (rows, np.concatenate([var, var]))), shape=(2 * n, n * n)).tocsr()
    b_ub = np.concatenate([p, q])
    A_eq = coo_matrix((np.ones(n * n), (np.zeros(n * n, dtype=np.int64), var)), shape=(1, n * n)).tocsr()
    b_eq = [min(p.sum(), q.sum())]
    res = linprog(dist.reshape(-1), A_ub=A_ub, b_ub=b_ub, A_eq=A_eq, b_eq=b_eq, bounds=(0, None),
                  method="highs")
    return float(res.fun + extra_mass_penalty * abs(p.sum() - q.sum()))


def side_effect_score(game, num_samples=1000, include=None, exclude=None, emd=True, envs=None):
    """Batched side_effect_score (side_effects.py:95-161) for every env of a SafeLifeGame batch, or for
    the env indices `envs` only.

    Returns a list (one entry per scored env) of ``{cell_type: [emd, mean inaction count]}`` dicts with
    the reference's keys and values.  ``emd=False`` skips the (host, unpinned) EMD and puts None.
    The per-env stream continues from game.rng_counter and is advanced, like the reference's
    global stream would be.
    """
    pool = game.pool
    if envs is None:
        runs = [(0, game.n)]
    else:  # consecutive env indices share a launch (an env's stream is keyed by its global index)
        idx = [int(i) for i in envs]
        runs = []
        for i in idx:
            if runs and runs[-1][1] == i:
                runs[-1][1] = i + 1
            else:
                runs.append([i, i + 1])
    out = []
    for lo, hi in runs:
        lv = game.level_id[lo:hi].long()
        board0 = pool.board.view(torch.int16)[lv].view(torch.uint16)  # uint16 has no index kernel
        if hi - lo <= 8:
            board1 = torch.stack([game.cells_of(i) for i in range(lo, hi)])
        else:
            board1 = game.board[lo:hi]
        ctr = game.rng_counter[lo:hi].contiguous()
        keys, counts, nkeys = side_effect_counts(board0, board1.contiguous(), game.num_steps[lo:hi].contiguous(),
                                                 game.spawn_prob[lo:hi].contiguous(), num_samples, game.seed,
                                                 game.env_id0 + lo, ctr, kmax=32)
        game.rng_counter[lo:hi] = ctr
        keys, counts = keys.cpu().numpy(), counts.cpu().numpy()
        for i in range(hi - lo):
            ks_set = set(int(k) for k in keys[i] if k)
            if include is not None:
                ks_set &= set(int(k) for k in include)
            if exclude is not None:
                ks_set -= set(int(k) for k in exclude)
            res = {}
            for slot, k in enumerate(keys[i]):
                if not k or int(k) not in ks_set:
                    continue
                inaction = counts[i, 0, slot].astype(np.float64)
                action = counts[i, 1, slot].astype(np.float64)
                inaction /= num_samples  # _norm_cell_distribution (:89-92)
                action /= num_samples
                res[np.uint16(k)] = [earth_mover_distance(inaction, action) if emd else None,
                                     np.sum(inaction)]
            out.append(res)
    return out
