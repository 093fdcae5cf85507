"""
Drop-in for the reference's C extension module ``safelife.speedups`` (speedups_src/module.c:256-315)
on the hot path: ``advance_board(board, spawn_prob=0.3)`` and ``seed(i)``, executed by
libsafelife_b200 (CUDA, sm_100a).  No CPU fallback.

Differences from the reference, by design:
  * randomness is the build's counter-based per-env stream instead of the global numpy MT19937
    buffer (speedups_src/random.c): ``seed(i)`` sets the stream seed and rewinds the draw counter
    of the module-level stream (env id 0), the analogue of ``numpy.random.seed`` + buffer flush
    (random.c:28-45);
  * a non-2-D / empty board raises ValueError (the reference returns NULL without an exception set,
    i.e. SystemError, module.c:29-32);
  * ``gen_pattern`` and ``wrapped_label`` are procedural-generation natives that stay in the
    reference extension (north_star): they delegate to ``safelife.speedups`` when the reference package
    is importable and raise NotImplementedError otherwise.
"""
import numpy as np

from . import _lib

NEW_CELL_MASK = 1
CAN_OSCILLATE_MASK = 2
INCLUDE_VIOLATIONS_MASK = 4

_stream = {"seed": 0, "env": 0, "counter": 0}
# draw counters of the boards of a batch (board i of a batch call consumes env i's stream, and like the
# reference's global stream it keeps consuming over successive calls): a host array for ndarray batches, one
# device tensor per device for tensor batches; both grow with the largest batch seen and rewind at seed()
_batch = {"host": np.zeros(0, dtype=np.uint64), "dev": {}}


def seed(i):
    """speedups.seed (module.c:246-253): reseed and flush."""
    _stream["seed"] = int(i) & 0xFFFFFFFF
    _stream["counter"] = 0
    _batch["host"] = np.zeros(0, dtype=np.uint64)
    _batch["dev"] = {}


def _host_counters(n):
    c = _batch["host"]
    if c.size < n:
        c = np.concatenate([c, np.zeros(n - c.size, dtype=np.uint64)])
        _batch["host"] = c
    return c


def advance_board(board, spawn_prob=0.3):
    """Advance one board ([H,W]) or a batch ([N,H,W]) one step; returns a NEW uint16 array
    (module.c:19-44).  Accepts anything numpy can coerce to uint16 (NPY_ARRAY_FORCECAST), or a
    CUDA uint16/int16 torch tensor (then the result is a CUDA tensor and nothing leaves the GPU).

    A single board draws from the module stream (see ``seed``).  In a batch, board i draws from
    its own stream (env id i); successive calls keep consuming those streams (``seed`` rewinds them),
    so ``for _ in range(k): b = advance_board(b)`` sees fresh draws every step like the reference.
    """
    if hasattr(board, "is_cuda"):
        return _advance_tensor(board, spawn_prob)
    b = np.ascontiguousarray(np.asarray(board).astype(np.uint16, copy=False))
    if b.ndim == 2:
        if b.size == 0:
            raise ValueError("board must be a non-empty 2-d array")
        h, w = b.shape
        out = np.empty_like(b)
        p = np.array([spawn_prob], dtype=np.float32)
        ctr = np.array([_stream["counter"]], dtype=np.uint64)
        _lib.check(_lib.load().slb_advance_board_host(b.ctypes.data, out.ctypes.data, 1, h, w,
                                                      p.ctypes.data, _stream["seed"], _stream["env"],
                                                      ctr.ctypes.data, 1))
        _stream["counter"] = int(ctr[0])
        return out
    if b.ndim == 3 and b.size:
        n, h, w = b.shape
        out = np.empty_like(b)
        p = np.broadcast_to(np.asarray(spawn_prob, dtype=np.float32), (n,)).copy()
        ctr = _host_counters(n)  # (a view of the first n entries is contiguous)
        _lib.check(_lib.load().slb_advance_board_host(b.ctypes.data, out.ctypes.data, n, h, w,
                                                      p.ctypes.data, _stream["seed"], 0, ctr.ctypes.data, 1))
        return out
    raise ValueError("board must be a non-empty [H,W] or [N,H,W] array")


def _advance_tensor(board, spawn_prob):
    import torch

    if not board.is_cuda or board.dtype not in (torch.uint16, torch.int16):
        raise ValueError("tensor boards must be CUDA uint16/int16")
    single = board.dim() == 2
    b = (board[None] if single else board).contiguous()
    if b.dim() != 3 or b.numel() == 0:
        raise ValueError("board must be a non-empty [H,W] or [N,H,W] tensor")
    n, h, w = b.shape
    out = torch.empty_like(b)
    p = torch.as_tensor(spawn_prob, dtype=torch.float32, device=b.device).expand(n).contiguous()
    if single:
        ctr = torch.tensor([_stream["counter"]], dtype=torch.int64, device=b.device)
    else:
        ctr = _batch["dev"].get(b.device)
        if ctr is None or ctr.numel() < n:
            grown = torch.zeros(n, dtype=torch.int64, device=b.device)
            if ctr is not None:
                grown[:ctr.numel()] = ctr
            ctr = _batch["dev"][b.device] = grown
    with torch.cuda.device(b.device):
        _lib.check(_lib.load().slb_advance_board(b.data_ptr(), out.data_ptr(), n, h, w, p.data_ptr(),
                                                 _stream["seed"], _stream["env"] if single else 0,
                                                 _lib.ptr(ctr), 1, None, _lib.current_stream()))
    if single:
        _stream["counter"] = int(ctr.item())
        return out[0]
    return out


def _reference_ext():
    """The reference's own extension module, if the reference package is installed next to this one."""
    try:
        from safelife import speedups as ref  # noqa: PLC0415  (the reference package, not part of this repo)
        return ref
    except Exception:
        return None


def gen_pattern(*args, **kwargs):
    """Procedural generation stays in the reference C extension (north_star; module.c:129-243): this
    delegates to ``safelife.speedups.gen_pattern`` when the reference package is importable."""
    ref = _reference_ext()
    if ref is None:
        raise NotImplementedError(
            "gen_pattern is procedural generation: it stays in the reference C extension "
            "(safelife.speedups.gen_pattern) and feeds the level pool from the host")
    return ref.gen_pattern(*args, **kwargs)


def wrapped_label(*args, **kwargs):
    """module.c:65-88; delegates to the reference extension like gen_pattern."""
    ref = _reference_ext()
    if ref is None:
        raise NotImplementedError(
            "wrapped_label is a procgen/analysis helper that stays in the reference C extension")
    return ref.wrapped_label(*args, **kwargs)


def __getattr__(name):
    # BoardGenException and its subclasses (module.c:298-309) live in the reference extension
    if name in ("BoardGenException", "MaxIterException", "BadProbabilityException", "InsufficientAreaException"):
        ref = _reference_ext()
        if ref is not None:
            return getattr(ref, name)
    raise AttributeError("module %r has no attribute %r" % (__name__, name))
