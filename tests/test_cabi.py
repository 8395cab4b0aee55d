"""CPU: the C-ABI library builds, loads and exports every symbol include/ccgpu.h declares
(no compute calls here - there is no GPU in the build container)."""
import ctypes
import os
import re

from consensuscruncher_b200 import build, engine

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_builds_and_exports_declared_symbols():
    path = build.build_lib()
    lib = ctypes.CDLL(path)
    header = open(os.path.join(ROOT, "include", "ccgpu.h")).read()
    declared = set(re.findall(r"\b(cc_[a-z_0-9]+)\s*\(", header))
    assert declared == set(engine.SYMBOLS)
    for sym in sorted(declared):
        assert hasattr(lib, sym), sym


def test_create_fails_loudly_without_a_gpu_or_reports_b200():
    lib = engine.load_library()
    params = engine.CcParams(0.7, 1)
    ctx = lib.cc_create(0, ctypes.byref(params))
    if not ctx:
        msg = lib.cc_last_error(None).decode()
        assert msg, "cc_create must explain why it failed"
    else:
        lib.cc_destroy(ctx)


def test_library_is_sm100a_only():
    import shutil
    import subprocess
    cuobjdump = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    out = subprocess.run([cuobjdump, "-lelf", build.LIB], capture_output=True, text=True).stdout
    assert "sm_100a" in out and "sm_90" not in out and "sm_80" not in out
