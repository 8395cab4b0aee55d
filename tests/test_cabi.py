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


def test_header_enums_and_record_sizes_agree_with_the_binding():
    """include/ccgpu.h is the contract: enum cc_array values, CC_ERR_* codes and the byte layout of the records a host
    hands over must be what engine.py / pack.py use."""
    from consensuscruncher_b200 import pack
    header = open(os.path.join(ROOT, "include", "ccgpu.h")).read()
    body = re.search(r"enum cc_array \{(.*?)\};", header, re.S).group(1)
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    values = {m.group(1): int(m.group(2)) for m in re.finditer(r"\b(CC_[A-Z_]+)\s*=\s*(\d+)", body)}
    name_of = {"fam_tmpl": "CC_FAM_TMPL", "read_fam": "CC_READ_FAM", "fam_hist": "CC_FAM_HIST", "stats": "CC_STATS"}
    for name, (what, _) in engine.ARRAYS.items():
        assert values[name_of.get(name, "CC_" + name.upper())] == what, name
    assert len(values) == len(engine.ARRAYS)
    errs = {m.group(1): int(m.group(2)) for m in re.finditer(r"\b(CC_ERR_[A-Z_]+)\s*=\s*(-\d+)", header)}
    assert errs["CC_ERR_BASE"] == engine.CC_ERR_BASE and errs["CC_ERR_DOMAIN"] == engine.CC_ERR_DOMAIN

    def c_sizeof(name):                        # the structs are plain 1 / 2 / 4-byte members without padding
        text = re.search(r"typedef struct \{([^}]*)\}\s*%s;" % name, header, re.S).group(1)
        text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
        width = {"uint32_t": 4, "int32_t": 4, "uint16_t": 2, "int16_t": 2, "uint8_t": 1}
        total = 0
        for typ, names in re.findall(r"\b(u?int(?:8|16|32)_t)\s+([^;]+);", text):
            for nm in names.split(","):
                dims = re.findall(r"\[(\d+)\]", nm)
                total += width[typ] * (int(dims[0]) if dims else 1)
        return total
    assert c_sizeof("cc_fields") == pack.FIELDS_DT.itemsize == 32
    assert c_sizeof("cc_pair") == pack.PAIR_DT.itemsize == 40
    assert c_sizeof("cc_pair16") == pack.PAIR16_DT.itemsize == 16
    assert c_sizeof("cc_pair_class") == pack.CLASS_DT.itemsize == 32
