"""
Import shim for running the UNMODIFIED reference package in the build container (oracle use only).

This file is ours (it replaces only the reference's safelife/__init__.py import guard).  It makes
``import safelife.<module>`` resolve reference modules from /root/reference/safelife (read-only,
never copied) and ``safelife.speedups`` from oracle/_ref/pyext (the reference C extension compiled
in place by oracle/Makefile).  /root/reference does not exist on the GPU box, so nothing that runs
there may import this.
"""
import os

_here = os.path.dirname(os.path.abspath(__file__))
_ref_root = os.environ.get("SAFELIFE_REFERENCE", "/root/reference")
_ref_pkg = os.path.join(_ref_root, "safelife")
_pyext = os.path.normpath(os.path.join(_here, "..", "..", "_ref", "pyext"))
if not os.path.isdir(_ref_pkg):
    raise ImportError("reference tree not present at %s" % _ref_pkg)
__path__ = [_here, _pyext, _ref_pkg]

from . import speedups  # noqa: E402,F401
