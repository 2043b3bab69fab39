"""The C-ABI library loads and exports every symbol include/reaspasol_gpu.h declares (no compute
calls here: this runs without a GPU)."""
import ctypes
import os
import re

from respasol_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_functions():
    text = open(os.path.join(ROOT, "include", "reaspasol_gpu.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    names = re.findall(r"^\s*(?:const\s+char\s*\*\s*|void\s*\*\s*|int\s+)(\w+)\s*\(", text, flags=re.M)
    return sorted(set(names))


def test_header_and_binding_tables_agree():
    assert header_functions() == sorted(_lib.GPU_SYMBOLS)


def test_library_exports_every_declared_symbol():
    assert os.path.exists(_lib.GPU_LIB), "libreaspasol_gpu.so not built (python -c 'import __graft_entry__ as g; g.build()')"
    lib = ctypes.CDLL(_lib.GPU_LIB)
    for name in header_functions():
        assert hasattr(lib, name), name
    for name in ("spmv_csr_s", "spmv_csr_d", "ilu0_factor_s", "ilu0_factor_d", "sptrsv_lu_s", "sptrsv_lu_d"):
        assert hasattr(lib, name)  # the six names BASELINE.json fixes


def test_status_strings_and_version():
    lib = _lib.gpu_lib()
    assert lib.rsg_version() >= 100
    assert lib.rsg_status_string(0) == b"ok"
    assert b"pivot" in lib.rsg_status_string(5)


def test_host_library_symbols():
    lib = _lib.host_lib()
    for name in _lib.HOST_SYMBOLS:
        assert hasattr(lib, name)


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "respasol_b200")
    for dirpath, _, files in os.walk(pkg):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                src = open(os.path.join(dirpath, fn), errors="ignore").read()
                assert "liboracle" not in src and "import orc" not in src and "orc_" not in src, fn


def _enum(name):
    """{identifier: value} of a typedef enum in the header (explicit values, or counting up from the previous one)"""
    text = open(os.path.join(ROOT, "include", "reaspasol_gpu.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    body = re.search(r"typedef enum \{([^}]*)\}\s*%s;" % name, text).group(1)
    out, nxt = {}, 0
    for item in filter(None, (s.strip() for s in body.split(","))):
        ident, _, val = item.partition("=")
        nxt = int(val.strip(), 0) if val.strip() else nxt
        out[ident.strip()] = nxt
        nxt += 1
    return out


def test_python_option_tables_mirror_the_header_enums():
    from respasol_b200 import gpu
    opts = _enum("rsg_option")
    assert {("RSG_OPT_" + k): v for k, v in gpu.OPT.items()} == {k: v for k, v in opts.items() if k != "RSG_OPT_COUNT_"}
    assert opts["RSG_OPT_COUNT_"] == max(gpu.OPT.values()) + 1
    algos = _enum("rsg_spmv_algo")
    for k, v in gpu.ALGO.items():
        assert algos["RSG_SPMV_" + k.upper()] == v, k
    status = _enum("rsg_status")
    lib = _lib.gpu_lib()
    for k, v in status.items():
        assert lib.rsg_status_string(v) != b"unknown status", k
    assert lib.rsg_status_string(max(status.values()) + 1) == b"unknown status"


def test_noalloc_scratch_size_stated_in_the_header():
    text = open(os.path.join(ROOT, "include", "reaspasol_gpu.h")).read()
    rows = 1 << int(re.search(r"#define RSG_NOALLOC_MAX_ROWS \(1 << (\d+)\)", text).group(1))
    tiles = 1 << int(re.search(r"#define RSG_NOALLOC_MAX_TILES \(1 << (\d+)\)", text).group(1))
    stated = int(re.search(r"#define RSG_NOALLOC_SCRATCH_BYTES (\d+)", text).group(1))
    desc = 16 * (tiles + 1 + 15)
    rp = (4 * (((rows + 1 + 3) & ~3) + 4) + 255) & ~255
    assert stated == desc + rp and stated < 8 << 20  # (the .cu static_asserts the same equality)
