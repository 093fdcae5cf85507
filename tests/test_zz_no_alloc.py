"""slb_emd_solve_static — the EMD solver for callers that cannot allocate (scratch in a static device array) —
against slb_emd_solve on the same inputs, through the C ABI, two calls in a row with different inputs (the
array keeps the previous call's contents).  Collected last on purpose (file name).
Passes alone in a fresh process and after the rest of the suite."""

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

H = W = 26
KMAX = 5
NS = 1000


def _maps(seed, n):
    """random count maps: sparse and dense differences, identical maps, one-sided maps, empty slots"""
    rng = np.random.RandomState(seed)
    keys = np.zeros((n, KMAX), np.uint16)
    counts = np.zeros((n, 2, KMAX, H, W), np.uint16)
    for i in range(n):
        for s in range(KMAX):
            kind = rng.randint(7)
            if kind == 0:
                continue
            keys[i, s] = 9 + s
            cells = int(rng.choice([1, 3, 10, 40, 250]))
            idx = rng.choice(H * W, cells, replace=False)
            a = np.zeros(H * W, np.uint16)
            a[idx] = rng.randint(1, NS + 1, cells)
            b = a.copy()
            if kind == 2:
                b[:] = 0
            elif kind > 2:
                move = rng.rand(cells) < 0.7
                b[idx[move]] = 0
                j = rng.choice(H * W, int(move.sum()))
                b[j] = rng.randint(1, NS + 1, j.size)
            counts[i, 0, s] = a.reshape(H, W)
            counts[i, 1, s] = b.reshape(H, W)
    return keys, counts


class _Problem:
    """everything both entry points need, allocated BEFORE the calls under test"""

    def __init__(self, lib, se, seed, n):
        from safelife_k2_b200 import _lib

        self.n = n
        keys, counts = _maps(seed, n)
        self.keys = torch.from_numpy(keys).cuda()
        self.counts = torch.from_numpy(counts).cuda()
        n_src = torch.empty((n, KMAX), dtype=torch.int32, device="cuda")
        n_snk = torch.empty_like(n_src)
        self.st = _lib.current_stream()
        _lib.check(lib.slb_emd_sizes(self.counts.data_ptr(), self.keys.data_ptr(), n, KMAX, H, W, NS,
                                     n_src.data_ptr(), n_snk.data_ptr(), self.st))
        v = (n_src + n_snk).flatten()
        self.order = torch.argsort(v, descending=True).to(torch.int32)
        self.n_problems = int((v > 0).sum().item())
        self.vmax = int(v.max().item())
        self.ft_stride = max(1, int((n_src.long() * n_snk.long()).max().item()))
        self.lut = torch.from_numpy(se.distance_table(H, W, "manhattan", 5.0)).cuda()
        self.scratch = torch.empty(64 * self.ft_stride, dtype=torch.float64, device="cuda")
        self.queue = torch.zeros(1, dtype=torch.int32, device="cuda")
        self.status = torch.zeros(2, dtype=torch.int32, device="cuda")
        self.emd = torch.full((2, n, KMAX), -7.0, dtype=torch.float64, device="cuda")

    def head(self):
        return (self.counts.data_ptr(), self.keys.data_ptr(), self.n, KMAX, H, W, NS, self.lut.data_ptr(), 1, 1,
                1.0, self.order.data_ptr(), self.n_problems, self.vmax, self.ft_stride)


def test_emd_solve_static_equals_emd_solve_twice_in_a_row():
    from safelife_k2_b200 import _lib
    from safelife_k2_b200 import side_effects as se

    lib = _lib.load()
    probs = [_Problem(lib, se, 11, 10), _Problem(lib, se, 12, 7)]   # different sizes: stale flows / queue matter
    assert all(p.n_problems > 8 and p.vmax > 100 for p in probs)
    torch.cuda.synchronize()
    free_before = None
    for call, p in enumerate(probs):
        if call == 1:
            free_before = torch.cuda.mem_get_info()[0]
        _lib.check(lib.slb_emd_solve_static(*p.head(), p.emd[1].data_ptr(), p.status[1:].data_ptr(), p.st))
        if call == 1:
            torch.cuda.synchronize()
            assert torch.cuda.mem_get_info()[0] == free_before      # nothing was allocated by the call
        _lib.check(lib.slb_emd_solve(*p.head(), p.scratch.data_ptr(), p.scratch.numel(), p.queue.data_ptr(),
                                     p.emd[0].data_ptr(), p.status[:1].data_ptr(), p.st))
        torch.cuda.synchronize()
        assert p.status.cpu().tolist() == [0, 0]
        want, got = p.emd[0].cpu().numpy().ravel(), p.emd[1].cpu().numpy().ravel()
        assert (np.abs(got - want) <= 1e-12 * np.maximum(1.0, np.abs(want))).all(), call
        touched = p.order[:p.n_problems].cpu().numpy()
        assert (want[touched] >= 0).all() and (want[touched] > 0).sum() > 8
        untouched = np.setdiff1d(np.arange(want.size), touched)
        assert (got[untouched] == -7.0).all() and (want[untouched] == -7.0).all()


def test_emd_solve_static_rejects_sizes_beyond_the_static_scratch():
    from safelife_k2_b200 import _lib
    from safelife_k2_b200 import side_effects as se

    lib = _lib.load()
    p = _Problem(lib, se, 13, 2)
    limit = 2097152 // 4                                            # SLB_EMD_STATIC_SCRATCH_DOUBLES / warps per CTA
    args = list(p.head())
    args[-1] = limit + 1
    rc = lib.slb_emd_solve_static(*args, p.emd[1].data_ptr(), p.status[1:].data_ptr(), p.st)
    assert rc == _lib.SLB_ERR_UNSUPPORTED
    msg = lib.slb_last_error().decode()
    assert "static scratch" in msg and str(limit) in msg, msg
    args[-1] = limit                                                # the limit itself is accepted
    _lib.check(lib.slb_emd_solve_static(*args, p.emd[1].data_ptr(), p.status[1:].data_ptr(), p.st))
    torch.cuda.synchronize()
    assert int(p.status[1].item()) == 0
