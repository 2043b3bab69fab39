"""numpy-side companions of the hot path: the CSR container (reference `struct CSR`,
ReadMatrixMarket/loadMatrixMarket.h:17-25), the Matrix-Market loader with the reference's
signature and quirks (loadMatrixMarket.cpp:47-253) and the synthetic SuiteSparse-shaped generators
(SURVEY.md 8d).  All heavy lifting is in librsg_host.so (C++); this file only moves arrays."""
import ctypes as C
from dataclasses import dataclass

import numpy as np

from ._lib import CCSR, host_lib

LOADER_BUGCOMPAT = 0
LOADER_EXPAND_SYMMETRIC = 1


@dataclass
class CSR:
    m: int
    n: int
    nnz: int                 # what the loader REPORTS (doubled on symmetric files)
    rowptr: np.ndarray       # int32, m+1
    colidx: np.ndarray       # int32
    values: np.ndarray       # float64 (the reference CSR is always fp64 on the host)
    base: int = 0
    is_symmetric: int = 0

    @property
    def stored(self):
        return int(self.rowptr[self.m] - self.rowptr[0])

    def spmv_bytes(self, precision="d"):
        """Algorithmic bytes of one product."""
        v, x = {"d": (8, 8), "s": (4, 4), "m": (4, 8)}[precision]
        return self.stored * (v + 4) + (self.m + 1) * 4 + self.m * x + self.n * x

    def rows(self, r0, r1):
        """Row slab [r0,r1) as an independent CSR (global column indices, rebased row pointers)."""
        lo, hi = int(self.rowptr[r0] - self.base), int(self.rowptr[r1] - self.base)
        rp = (self.rowptr[r0:r1 + 1] - self.rowptr[r0] + self.base).astype(np.int32)
        return CSR(r1 - r0, self.n, hi - lo, rp, self.colidx[lo:hi].copy(), self.values[lo:hi].copy(), self.base)


def _take(c: CCSR, base: int) -> CSR:
    lib = host_lib()
    m = c.m
    stored = lib.rsg_csr_stored_nnz(C.byref(c))
    rowptr = np.ctypeslib.as_array(c.rowptr, shape=(m + 1,)).copy()
    colidx = np.ctypeslib.as_array(c.colidx, shape=(max(stored, 1),))[:stored].copy()
    values = np.ctypeslib.as_array(c.values, shape=(max(stored, 1),))[:stored].copy()
    out = CSR(m, c.n, c.nnz, rowptr.astype(np.int32), colidx.astype(np.int32), values, base, c.isSymmetric)
    lib.rsg_free_csr(C.byref(c))
    return out


def _as_c(a: CSR) -> CCSR:
    c = CCSR()
    c.isSymmetric, c.m, c.n, c.nnz = a.is_symmetric, a.m, a.n, a.stored
    c.rowptr = a.rowptr.ctypes.data_as(C.POINTER(C.c_int))
    c.colidx = a.colidx.ctypes.data_as(C.POINTER(C.c_int))
    c.values = a.values.ctypes.data_as(C.POINTER(C.c_double))
    return c


def load_matrix_market(path, output_base=0, transpose=0, flags=LOADER_BUGCOMPAT) -> CSR:
    """loadMatrixMarket(file, &A, outputBase, transpose): 1 on success in C; raises here on failure."""
    c = CCSR()
    ok = host_lib().rsg_load_matrix_market(str(path).encode(), C.byref(c), output_base, transpose, flags)
    if ok != 1:
        raise ValueError(f"loadMatrixMarket failed for {path}")
    return _take(c, output_base)


def write_matrix_market(path, a: CSR):
    if host_lib().rsg_write_matrix_market(str(path).encode(), C.byref(_as_c(a)), a.base) != 1:
        raise OSError(f"cannot write {path}")


def _gen(fn, *args) -> CSR:
    c = CCSR()
    if fn(*args, C.byref(c)) != 1:
        raise ValueError("generator rejected its arguments")
    return _take(c, 0)


def gen_laplace2d(N) -> CSR:
    """C1: 5-point Laplacian on N x N (ecology2-shaped); nnz = 5N^2-4N."""
    return _gen(host_lib().rsg_gen_laplace2d, N)


def gen_stencil27(nx, ny=None, nz=None, z0=0, z1=None, seed=0) -> CSR:
    """C2/C5: 27-point stencil; rows of planes [z0,z1) only, global columns."""
    ny = nx if ny is None else ny
    nz = nx if nz is None else nz
    z1 = nz if z1 is None else z1
    return _gen(host_lib().rsg_gen_stencil27, nx, ny, nz, z0, z1, seed)


def gen_femlike(N, p=0.58, seed=3) -> CSR:
    """C3: FEM-like, strictly diagonally dominant (2cubes_sphere-shaped) for ILU0."""
    return _gen(host_lib().rsg_gen_femlike, N, p, seed)


def gen_powerlaw(m, alpha=2.2, maxlen=64, ndense=4, dense_len=100000, seed=4) -> CSR:
    """C4: power-law row lengths plus a few dense rows/columns (ASIC_320ks-shaped)."""
    return _gen(host_lib().rsg_gen_powerlaw, m, alpha, maxlen, ndense, dense_len, seed)


def gen_banded(m, per_row, half_band, seed=7) -> CSR:
    return _gen(host_lib().rsg_gen_banded, m, per_row, half_band, seed)


def dlarnv_u01(n) -> np.ndarray:
    """x = LAPACKE_dlarnv(1, {0,0,0,1}, n) restated (test_spmv.c:75-76)."""
    x = np.empty(n, dtype=np.float64)
    host_lib().rsg_dlarnv_u01(n, x.ctypes.data_as(C.POINTER(C.c_double)))
    return x
