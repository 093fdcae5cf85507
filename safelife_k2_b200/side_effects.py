"""
Side-effect score — host-side mirror of the reference's ``safelife/side_effects.py``.

The expensive part (side_effects.py:131-141: catch-up of the inaction board, 2 x num_samples
paired advance_board steps, per-cell-type count maps of _add_cell_distribution :59-86) runs on the
GPU for all finished episodes at once (slb_side_effect_counts).  The host then does what the
reference does with the maps: divide by n (:89-92), np.sum, and the earth-mover distance.

EMD: the reference calls third-party ``pyemd.emd`` (pyemd 0.5.1, not vendored, not installed
here) at side_effects.py:56, i.e. Pele & Werman's EMD-hat, ``emd_hat_gd_metric``: min(a_i, b_i) is cancelled
inside every bin, the rest is an exact min-cost transport, the surplus pays extra_mass_penalty per unit.
``side_effect_score`` computes that for every (episode, cell type) pair of a batch on the GPU
(``emd_batch`` -> slb_emd_sizes / slb_emd_solve: one warp per transportation problem, successive shortest
augmenting paths).  ``earth_mover_distance(a, b)`` keeps the reference's signature for a single pair of float
density maps on the host (pyemd if importable, else scipy's LP of the same reduced problem).  That one number is
"parity unpinned" against pyemd; everything before it is bit-exact.
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


def distance_table(h, w, metric="manhattan", tanh_scale=5.0):
    """Ground distance of two cells whose coordinates differ by (|dy|, |dx|) after the wrap, float64 [H,W],
    with the reference's own numpy expressions (side_effects.py:50-55) so that the values are the ones it uses."""
    dy, dx = np.meshgrid(np.arange(h), np.arange(w), indexing="ij")
    if metric == "manhattan":
        dist = (np.abs(dx) + np.abs(dy)).astype(float)
    else:
        dist = np.sqrt(dx * dx + dy * dy)
    if tanh_scale > 0:
        dist = np.tanh(dist / tanh_scale)
    return np.ascontiguousarray(dist, dtype=np.float64)


def emd_batch(keys, counts, num_samples, metric="manhattan", wrap_x=True, wrap_y=True, tanh_scale=5.0,
              extra_mass_penalty=1.0):
    """earth_mover_distance(inaction density, action density) for every key slot of every episode, on the GPU.
    keys uint16 [N,kmax] (0 = empty), counts uint16 [N,2,kmax,H,W] (the outputs of side_effect_counts).
    -> float64 cuda tensor [N,kmax] (0 for empty slots)."""
    n, _, kmax, h, w = counts.shape
    dev = counts.device
    lib = _lib.load()
    keys = keys.contiguous()
    counts = counts.contiguous()
    emd = torch.zeros((n, kmax), dtype=torch.float64, device=dev)
    if n == 0:
        return emd
    n_src = torch.empty((n, kmax), dtype=torch.int32, device=dev)
    n_snk = torch.empty_like(n_src)
    with torch.cuda.device(dev):
        st = _lib.current_stream()
        _lib.check(lib.slb_emd_sizes(counts.data_ptr(), keys.data_ptr(), n, kmax, h, w, int(num_samples),
                                     n_src.data_ptr(), n_snk.data_ptr(), st))
        v = (n_src + n_snk).flatten()
        order = torch.argsort(v, descending=True).to(torch.int32)
        n_problems = int((v > 0).sum().item())        # (one host sync: sizes the solver)
        if n_problems == 0:
            return emd
        vmax = int(v.max().item())
        ft_stride = max(1, int((n_src.to(torch.int64) * n_snk.to(torch.int64)).max().item()))
        if extra_mass_penalty == -1:                   # pyemd's default: the largest ground distance
            extra_mass_penalty = float(distance_table(h, w, metric, tanh_scale).max())
        lut = torch.from_numpy(distance_table(h, w, metric, tanh_scale)).to(dev)
        sms = torch.cuda.get_device_properties(dev).multi_processor_count
        warps = min(4 * ((n_problems + 3) // 4), sms * 16)
        budget = 2 << 30                               # bytes of flow-matrix scratch at most
        warps = max(4, min(warps, (budget // 8 // ft_stride) // 4 * 4))
        scratch = torch.empty(warps * ft_stride, dtype=torch.float64, device=dev)
        queue = torch.zeros(1, dtype=torch.int32, device=dev)
        status = torch.zeros(1, dtype=torch.int32, device=dev)
        _lib.check(lib.slb_emd_solve(counts.data_ptr(), keys.data_ptr(), n, kmax, h, w, int(num_samples),
                                     lut.data_ptr(), int(bool(wrap_x)), int(bool(wrap_y)), float(extra_mass_penalty),
                                     order.data_ptr(), n_problems, vmax, ft_stride, scratch.data_ptr(),
                                     scratch.numel(), queue.data_ptr(), emd.data_ptr(), status.data_ptr(), st))
        if int(status.item()):
            raise _lib.SlbError(_lib.SLB_ERR_CUDA, "slb_emd_solve: a transportation problem did not converge")
    return emd


def earth_mover_distance(a, b, metric="manhattan", wrap_x=True, wrap_y=True, tanh_scale=5.0,
                         extra_mass_penalty=1.0):
    """side_effects.py:12-56 for one pair of float density maps, on the host (reference signature).  UNPINNED
    against pyemd when pyemd is absent (see the module docstring); side_effect_score itself uses emd_batch."""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    height, width = a.shape
    diff = np.abs(a - b)
    ys, xs = np.nonzero(diff > 1e-3 * diff.max())      # the cells that count as changed (:39-40)
    if ys.size == 0:
        return 0.0
    # coordinate differences bin i -> bin j; the reference wraps min(d, size - d) of the SIGNED difference
    # (:44-49), so only positive differences wrap and the ground distance is not symmetric
    ddx = xs[:, None] - xs[None, :]
    ddy = ys[:, None] - ys[None, :]
    if wrap_x:
        ddx = np.minimum(ddx, width - ddx)
    if wrap_y:
        ddy = np.minimum(ddy, height - ddy)
    dist = distance_table(height, width, metric, tanh_scale)[np.abs(ddy), np.abs(ddx)]
    p, q = a[ys, xs], b[ys, xs]
    try:
        import pyemd
        return pyemd.emd(p, q, dist, extra_mass_penalty)
    except ImportError:
        return _emd_lp(p, q, dist, extra_mass_penalty)


def _emd_lp(p, q, dist, extra_mass_penalty):
    """EMD-hat as pyemd.emd computes it (emd_hat_gd_metric), with scipy's LP for the transport: the mass
    min(p_i, q_i) is cancelled inside every bin, the rest flows from the bins that only supply to the bins that
    only demand at minimum cost, and the surplus pays extra_mass_penalty per unit."""
    from scipy.optimize import linprog
    from scipy.sparse import coo_matrix

    p = np.asarray(p, dtype=np.float64)
    q = np.asarray(q, dtype=np.float64)
    if extra_mass_penalty == -1:
        extra_mass_penalty = dist.max()
    both = np.minimum(p, q)
    pr, qr = p - both, q - both
    src, dst = np.nonzero(pr > 0)[0], np.nonzero(qr > 0)[0]
    surplus = (max(p.sum(), q.sum()) - min(p.sum(), q.sum())) * extra_mass_penalty
    n, m = src.size, dst.size
    if n == 0 or m == 0:
        return float(surplus)
    # flow f[i, j] >= 0 (variable i * m + j): row sums <= supply, column sums <= demand, total = the smaller
    # side; the constraint matrix has 2 n m ones — kept sparse
    var = np.arange(n * m)
    rows = np.concatenate([var // m, n + var % m])
    A_ub = coo_matrix((np.ones(2 * n * m), (rows, np.concatenate([var, var]))), shape=(n + m, n * m)).tocsr()
    A_eq = coo_matrix((np.ones(n * m), (np.zeros(n * m, dtype=np.int64), var)), shape=(1, n * m)).tocsr()
    res = linprog(dist[np.ix_(src, dst)].reshape(-1), A_ub=A_ub, b_ub=np.concatenate([pr[src], qr[dst]]),
                  A_eq=A_eq, b_eq=[min(pr[src].sum(), qr[dst].sum())], bounds=(0, None), method="highs")
    return float(res.fun + surplus)


def side_effect_score(game, num_samples=1000, include=None, exclude=None, emd=True, envs=None):
    """Batched side_effect_score (side_effects.py:95-161) for every env of a SafeLifeGame batch, or for
    the env indices `envs` only.

    Returns a list (one entry per scored env) of ``{cell_type: [emd, mean inaction count]}`` dicts with
    the reference's keys and values.  The EMDs of the whole batch are computed on the GPU (emd_batch; "unpinned"
    against pyemd, see the module docstring); ``emd=False`` skips them and puts None.
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
        emds = emd_batch(keys, counts, num_samples).cpu().numpy() if emd else None
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
                res[np.uint16(k)] = [float(emds[i, slot]) if emd else None, np.sum(inaction)]
            out.append(res)
    return out
