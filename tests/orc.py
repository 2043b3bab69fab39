"""ctypes doorway onto the CPU oracle (oracle/liboracle.so) and, when present, the reference's own
Matrix-Market loader compiled into oracle/_ref.  TEST INFRASTRUCTURE: imported only by tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
LIB = os.path.join(ORACLE_DIR, "liboracle.so")
REF_LIB = os.path.join(ORACLE_DIR, "_ref", "libloadmatrix_ref.so")

_lib = None
_ref = None
_ip, _dp, _fp = C.POINTER(C.c_int), C.POINTER(C.c_double), C.POINTER(C.c_float)


def build():
    subprocess.run(["make", "-C", ORACLE_DIR, "all"], check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB):
            build()
        _lib = C.CDLL(LIB)
        _lib.orc_max_threads.restype = C.c_int
    return _lib


def _i(a):
    return a.ctypes.data_as(_ip)


def _v(a):
    return a.ctypes.data_as(C.c_void_p)


def dlarnv(n):
    x = np.empty(n, dtype=np.float64)
    lib().orc_dlarnv_u01(C.c_int64(n), _v(x))
    return x


def spmv(a, x, precision="d", alpha=1.0, beta=0.0, y=None, threads=0):
    """precision 'd' | 's' | 'm'; values/x are cast as the reference harness does (spmv.cu:63-67)."""
    vt = np.float64 if precision == "d" else np.float32
    xt = np.float32 if precision == "s" else np.float64
    vals = np.ascontiguousarray(a.values, dtype=vt)
    x = np.ascontiguousarray(x, dtype=xt)
    y = np.zeros(a.m, dtype=xt) if y is None else np.array(y, dtype=xt)
    name = {"d": "orc_spmv_csr_d", "s": "orc_spmv_csr_s", "m": "orc_spmv_csr_mixed"}[precision]
    sc = C.c_float if precision == "s" else C.c_double
    if threads:
        getattr(lib(), name + "_omp")(C.c_int(threads), C.c_int(a.m), C.c_int(a.base), _i(a.rowptr), _i(a.colidx), _v(vals),
                                      sc(alpha), _v(x), sc(beta), _v(y))
    else:
        getattr(lib(), name)(C.c_int(a.m), C.c_int(a.base), _i(a.rowptr), _i(a.colidx), _v(vals), sc(alpha), _v(x),
                             sc(beta), _v(y))
    return y


def ilu0(a, dtype=np.float64):
    """returns (status, pivot_row, lu_values)"""
    vals = np.array(a.values, dtype=dtype)
    piv = C.c_int(-1)
    fn = lib().orc_ilu0_d if dtype == np.float64 else lib().orc_ilu0_s
    fn.restype = C.c_int
    st = fn(C.c_int(a.m), C.c_int(a.base), _i(a.rowptr), _i(a.colidx), _v(vals), C.byref(piv))
    return st, piv.value, vals


def sptrsv_lu(a, lu, b):
    """returns (status, z, y)"""
    lu = np.ascontiguousarray(lu)
    b = np.ascontiguousarray(b, dtype=lu.dtype)
    z = np.zeros(a.m, dtype=lu.dtype)
    y = np.zeros(a.m, dtype=lu.dtype)
    fn = lib().orc_sptrsv_lu_d if lu.dtype == np.float64 else lib().orc_sptrsv_lu_s
    fn.restype = C.c_int
    st = fn(C.c_int(a.m), C.c_int(a.base), _i(a.rowptr), _i(a.colidx), _v(lu), _v(b), _v(z), _v(y))
    return st, z, y


def levels(a, which):
    """returns (lev, level_ptr, level_rows)"""
    n = a.m
    lev = np.zeros(max(n, 1), dtype=np.int32)
    nl = C.c_int(0)
    lp = np.zeros(n + 2, dtype=np.int32)
    rows = np.zeros(max(n, 1), dtype=np.int32)
    lib().orc_levels(C.c_int(n), C.c_int(a.base), _i(a.rowptr), _i(a.colidx), C.c_int(which), _i(lev), C.byref(nl),
                     _i(lp), _i(rows))
    return lev[:n], lp[:nl.value + 1].copy(), rows[:n]


def row_partition(a, P, balanced=True):
    out = np.zeros(P + 1, dtype=np.int32)
    lib().orc_row_partition(C.c_int(a.m), C.c_int(a.base), _i(a.rowptr), C.c_int(P), C.c_int(int(balanced)), _i(out))
    return out


def max_threads():
    return lib().orc_max_threads()


# ---- the reference's own loader (only where oracle/_ref was built) ---------------------------
class _RefCSR(C.Structure):
    _fields_ = [("isSymmetric", C.c_int), ("m", C.c_int), ("n", C.c_int), ("nnz", C.c_int),
                ("rowptr", _ip), ("colidx", _ip), ("values", _dp)]


def have_ref():
    return os.path.exists(REF_LIB)


def ref_load(path, base=0, transpose=0):
    """dict with the reference loader's output (entries actually stored: rowptr[m]-rowptr[0])."""
    global _ref
    if _ref is None:
        _ref = C.CDLL(REF_LIB)
        _ref.ref_loadMatrixMarket.restype = C.c_int
    c = _RefCSR()
    ok = _ref.ref_loadMatrixMarket(str(path).encode(), C.byref(c), base, transpose)
    if ok != 1:
        return {"ok": int(ok)}
    m = c.m
    rowptr = np.ctypeslib.as_array(c.rowptr, shape=(m + 1,)).copy()
    stored = int(rowptr[m] - rowptr[0])
    colidx = np.ctypeslib.as_array(c.colidx, shape=(max(stored, 1),))[:stored].copy()
    values = np.ctypeslib.as_array(c.values, shape=(max(stored, 1),))[:stored].copy()
    out = {"ok": 1, "m": m, "n": c.n, "nnz": c.nnz, "isSymmetric": c.isSymmetric, "rowptr": rowptr, "colidx": colidx,
           "values": values}
    _ref.ref_free_csr(C.byref(c))
    return out


# ---- the reference's arithmetic library on the GPU box: cuSPARSE (oracle/libcusparse_ref.so) ----
CUSPARSE_LIB = os.path.join(ORACLE_DIR, "libcusparse_ref.so")
_cus = None


def cusparse():
    """The reference harness's cuSPARSE call groups (oracle/cusparse_ref.cu); needs a GPU to call."""
    global _cus
    if _cus is None:
        if not os.path.exists(CUSPARSE_LIB):
            build()
        _cus = C.CDLL(CUSPARSE_LIB)
        _cus.cref_spmv.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int] + [C.c_void_p] * 5 + \
            [C.c_int, C.c_int, C.c_void_p, C.c_size_t, C.c_void_p]
        _cus.cref_spmv_t.argtypes = _cus.cref_spmv.argtypes
        _cus.cref_ilu0.argtypes = [C.c_int] * 4 + [C.c_void_p] * 5
        _cus.cref_spsv_lu.argtypes = [C.c_int] * 3 + [C.c_void_p] * 6 + [C.c_int, C.c_void_p]
    return _cus


def _cus_prec(t):
    import torch
    return {torch.float64: 0, torch.float32: 1}[t.dtype]


def cusparse_spmv(A, x, y=None, alg=0, warm=0, reps=0, flush=None, transpose=False):
    """cusparseSpMV on a DeviceCSR ('d' or 's', base 0) -> (y, [mean, median, min] us or None); transpose: y = A^T x."""
    import torch
    assert A.base == 0 and A.values.dtype == x.dtype
    y = torch.empty(A.n if transpose else A.m, dtype=x.dtype, device=x.device) if y is None else y
    us = (C.c_float * 3)()
    fn = cusparse().cref_spmv_t if transpose else cusparse().cref_spmv
    st = fn(_cus_prec(x), alg, A.m, A.n, A.stored, A.rowptr.data_ptr(), A.colidx.data_ptr(),
                              A.values.data_ptr(), x.data_ptr(), y.data_ptr(), warm, reps,
                              flush.data_ptr() if flush is not None else None,
                              flush.numel() * flush.element_size() if flush is not None else 0, us)
    assert st == 0, "cref_spmv status %d" % st
    return y, (list(us) if reps else None)


def cusparse_ilu0(A, values, level=0):
    """csrilu02 in place on `values` -> (structural pivot, numeric pivot, analysis ms, numeric ms)."""
    piv = (C.c_int * 2)()
    ms = (C.c_float * 2)()
    st = cusparse().cref_ilu0(_cus_prec(values), level, A.m, A.stored, A.rowptr.data_ptr(), A.colidx.data_ptr(),
                              values.data_ptr(), piv, ms)
    assert st == 0, "cref_ilu0 status %d" % st
    return piv[0], piv[1], ms[0], ms[1]


def cusparse_spsv_lu(A, lu, b, reps=1):
    """cusparseSpSV L (unit lower) then U (non-unit upper) -> (z, y, analysis ms, solve ms)."""
    import torch
    z, y = torch.empty_like(b), torch.empty_like(b)
    ms = (C.c_float * 2)()
    st = cusparse().cref_spsv_lu(_cus_prec(lu), A.m, A.stored, A.rowptr.data_ptr(), A.colidx.data_ptr(), lu.data_ptr(),
                                 b.data_ptr(), z.data_ptr(), y.data_ptr(), reps, ms)
    assert st == 0, "cref_spsv_lu status %d" % st
    return z, y, ms[0], ms[1]


def bicgstab(a, b, lu=None, x0=None, max_iters=1000, rtol=1e-8):
    """oracle BiCGStab (oracle.c: orc_bicgstab_d), same conventions as pcg"""
    return pcg(a, b, lu, x0, max_iters, rtol, _fn="orc_bicgstab_d")


def gmres(a, b, lu=None, x0=None, max_iters=1000, rtol=1e-8):
    """oracle GMRES(30) (oracle.c: orc_gmres_d), same conventions as pcg; an iteration is one Arnoldi step"""
    return pcg(a, b, lu, x0, max_iters, rtol, _fn="orc_gmres_d")


def pcg(a, b, lu=None, x0=None, max_iters=1000, rtol=1e-8, _fn="orc_pcg_d"):
    """oracle PCG (oracle.c: orc_pcg_d) -> (x, iterations, relres); lu: None | fp64 | fp32 ILU(0) factor values"""
    x = np.zeros(a.m) if x0 is None else np.array(x0, dtype=np.float64)
    rel = C.c_double(0.0)
    prec = 0 if lu is None else (ord("d") if lu.dtype == np.float64 else ord("s"))
    fn = getattr(lib(), _fn)
    fn.restype = C.c_int
    vals = np.ascontiguousarray(a.values, dtype=np.float64)
    b = np.ascontiguousarray(b, dtype=np.float64)
    it = fn(C.c_int(a.m), C.c_int(a.base), _i(a.rowptr), _i(a.colidx), _v(vals), _v(lu) if lu is not None else None,
            C.c_int(prec), _v(b), _v(x), C.c_int(max_iters), C.c_double(rtol), C.byref(rel))
    return x, it, rel.value
