"""CPU: the C-ABI library loads and exports every symbol include/safelife_b200.h declares; the
argument checks that run before any CUDA call behave; the product never imports the oracle."""
import ctypes
import os
import re

from tests.conftest import ROOT

HEADER = os.path.join(ROOT, "include", "safelife_b200.h")


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(slb_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from safelife_k2_b200 import _lib

    names = declared_symbols()
    assert len(names) >= 16
    lib = ctypes.CDLL(_lib.LIB_PATH)
    for n in names:
        assert hasattr(lib, n), "libsafelife_b200.so does not export %s" % n
    assert sorted(_lib.SYMBOLS) == names, "python binding table out of sync with the header"
    assert _lib.load().slb_abi_version() == 5


def test_env_state_layout_matches_header():
    from safelife_k2_b200.safelife_game import ENV_STATE_DTYPE
    from oracle.slo import ENV_STATE_DTYPE as ORACLE_DTYPE

    src = open(HEADER).read()
    body = src[src.index("typedef struct {", src.index("Per-env scalar state")):src.index("} slb_env_state;")]
    fields = []
    for ctype, names in re.findall(r"\b(int32_t|int64_t|uint64_t|uint32_t|double)\s+([a-z_, ]+);", body):
        fields += [n.strip() for n in names.split(",")]
    assert fields == list(ENV_STATE_DTYPE.names) == list(ORACLE_DTYPE.names)
    assert ENV_STATE_DTYPE.itemsize == 80


def test_argument_checks_without_gpu():
    from safelife_k2_b200 import _lib

    L = _lib.load()
    one = ctypes.create_string_buffer(64)
    a = ctypes.addressof(one)
    assert L.slb_advance_board(a, a + 32, 1, 2, 5, a, 0, 0, None, 1, None, None) == _lib.SLB_ERR_INVALID
    assert b"3x3" in L.slb_last_error()
    assert L.slb_advance_board(a, a, 1, 4, 4, a, 0, 0, None, 1, None, None) == _lib.SLB_ERR_INVALID
    assert L.slb_advance_board(a, a + 32, 1, 4, 4, a, 0, 0, None, 0, None, None) == _lib.SLB_ERR_INVALID
    assert L.slb_advance_board(None, a, 1, 4, 4, a, 0, 0, None, 1, None, None) == _lib.SLB_ERR_INVALID
    assert L.slb_pool_prepare(None, None) == _lib.SLB_ERR_INVALID
    assert L.slb_side_effect_counts(a, a, 1, 4, 4, a, 10, a, 0, 0, None, 1000, a, a, a, None, None) \
        == _lib.SLB_ERR_INVALID
    # the bound steppers and the measurement aid reject bad arguments before touching the device
    out = ctypes.c_void_p()
    assert L.slb_stepper_create(None, 1, a, a, a, 1000, 0, 0, 0, None, ctypes.byref(out)) == _lib.SLB_ERR_INVALID
    assert L.slb_stepper_step(None, a, a, a) == _lib.SLB_ERR_INVALID
    assert L.slb_stepper_wait(None) == _lib.SLB_ERR_INVALID
    assert L.slb_dev_stepper_create(None, 1, a, a, a, 1000, 0, 0, 0, a, 15, 15, 1, None, 0, None, ctypes.byref(out)) \
        == _lib.SLB_ERR_INVALID
    assert L.slb_dev_stepper_step(None, a, a, a, a, None) == _lib.SLB_ERR_INVALID
    assert L.slb_microbench_int(2, 1, 1, a, None, None) == _lib.SLB_ERR_INVALID
    assert L.slb_microbench_int(0, 0, 1, a, None, None) == _lib.SLB_ERR_INVALID


def test_product_does_not_touch_the_oracle():
    pkg = os.path.join(ROOT, "safelife_k2_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in text.replace("parity oracle", ""), f
                assert "slo_" not in text, f


def test_procgen_natives_delegate_to_the_reference_extension(monkeypatch):
    """speedups.gen_pattern / wrapped_label / the BoardGen exceptions stay in the reference extension
    (SURVEY 8(b)): delegated when `safelife.speedups` is importable, NotImplementedError otherwise."""
    import sys
    import types

    import pytest

    from safelife_k2_b200 import speedups

    monkeypatch.setitem(sys.modules, "safelife", None)  # not importable
    with pytest.raises(NotImplementedError):
        speedups.gen_pattern(None, None, 1)
    with pytest.raises(AttributeError):
        speedups.BoardGenException
    pkg, ext = types.ModuleType("safelife"), types.ModuleType("safelife.speedups")
    ext.gen_pattern = lambda *a, **k: ("gen", a, k)
    ext.wrapped_label = lambda data: ("label", data)
    ext.BoardGenException = type("BoardGenException", (Exception,), {})
    pkg.speedups = ext
    monkeypatch.setitem(sys.modules, "safelife", pkg)
    monkeypatch.setitem(sys.modules, "safelife.speedups", ext)
    assert speedups.gen_pattern(1, 2, period=3) == ("gen", (1, 2), {"period": 3})
    assert speedups.wrapped_label(7) == ("label", 7)
    assert speedups.BoardGenException is ext.BoardGenException


def test_fast_binding_of_the_host_step_calls():
    """safelife_k2_b200/_slbfast (csrc/slb_pyfast.c, built by build()): the CPython binding of
    slb_stepper_step / slb_stepper_wait lands in the same library instance as the ctypes handle."""
    from safelife_k2_b200 import _lib, _slbfast

    L = _lib.load()
    assert _slbfast.step(0, 0, 0, 0) == _lib.SLB_ERR_INVALID
    assert b"stepper is NULL" in L.slb_last_error()
    assert _slbfast.wait(0) == _lib.SLB_ERR_INVALID
    assert _slbfast.step(None, None, None, None) == _lib.SLB_ERR_INVALID


def test_every_product_module_imports_without_a_gpu():
    """host-side modules must at least import (and the benches compile) on a box without a GPU"""
    import importlib
    import py_compile

    for m in ("_lib", "safelife_game", "safelife_env", "env_wrappers", "side_effects", "speedups", "benchmarking",
              "distributed"):
        importlib.import_module("safelife_k2_b200." + m)
    for f in ("bench.py", "bench_configs.py", "__graft_entry__.py", "scripts/sanitize_paths.py"):
        py_compile.compile(os.path.join(ROOT, f), doraise=True)


def test_bench_reference_arm_line():
    """`bench.py --impl reference` (the CPU arm the driver runs beside the GPU arm): one JSON line with the contract's
    keys, timed on the oracle's CPU path (no GPU needed).  SLB_REF_SECONDS keeps the sample short here."""
    import json
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, SLB_REF_SECONDS="1.5")
    out = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--impl", "reference", "--steps", "3",
                          "--warmup", "1"], capture_output=True, text=True, timeout=300, env=env, cwd=root)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "env_steps_per_s" and d["unit"] == "env-steps/s"
    assert d["higher_is_better"] is True and d["steps"] == 3 and d["value"] > 0
    assert d["cpu_baseline"]["cores"] >= 1 and d["cpu_baseline"]["value"] == d["value"]
    assert d["e2e"] == {"value": d["value"], "unit": "env-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["config"]["envs_per_gpu"] == 4096 and d["config"]["board"] == [26, 26]
