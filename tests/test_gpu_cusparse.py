"""Parity against the library the reference itself computes with: cuSPARSE, called live on the GPU
box through oracle/libcusparse_ref.so (the reference harness's call groups, GPU/spmv.cu:132-186,
GPU/ilu0.cu:166-310, restated for cuSPARSE 12.x), on the SAME device arrays libreaspasol_gpu reads.

cuSPARSE fixes no summation order, so nothing here is bit-exact; tolerances are the ones north_star /
SURVEY 8c state:
  fp64 SpMV     ||y - y_cusparse||inf <= 1e-12 ||y_cusparse||inf
  fp32 SpMV     per row |d_i| <= 2 (len_i + 2) 2^-24 sum_k |a_ik x_k|        (both sides round)
  ILU0 factors  componentwise |d_ij| <= c eps (|L||U|)_ij, c = 4 (longest row + 2)
  L/U solves    componentwise against c eps (|L^-1| ... ) is not computable cheaply: normwise
                ||y - y_cusparse||inf <= 1e-10 (fp64) / 2e-4 (fp32) ||y||inf on diagonally dominant input
Pivot reports (structural / numerical zero, row index) must be identical integers."""
import os

import numpy as np
import pytest
import torch

import orc
from respasol_b200 import host
from respasol_b200.gpu import DeviceCSR

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")

SPMV_MATS = {
    "lap2d_300": lambda: host.gen_laplace2d(300),
    "stencil27_40": lambda: host.gen_stencil27(40, seed=2),
    "femlike_24": lambda: host.gen_femlike(24),
    "powerlaw_20k": lambda: host.gen_powerlaw(20000, 2.2, 64, 3, 6000, 4),
    "bcspwr01": lambda: host.load_matrix_market(os.path.join(GOLD, "bcspwr01.mtx")),
}


def _row_abs_sums(a, x):
    rp = a.rowptr.astype(np.int64) - a.base
    t = np.abs(a.values.astype(np.float64) * x.astype(np.float64)[a.colidx.astype(np.int64) - a.base])
    s = np.concatenate([[0.0], np.cumsum(t)])
    return s[rp[1:]] - s[rp[:-1]], np.diff(rp)


@pytest.mark.parametrize("name", sorted(SPMV_MATS))
@pytest.mark.parametrize("alg", [0, 1, 2])
def test_spmv_fp64_matches_cusparse(handle, name, alg):
    a = SPMV_MATS[name]()
    A = DeviceCSR(a, "d")
    x = torch.from_numpy(orc.dlarnv(a.n) - 0.5).cuda()
    y = handle.spmv(A, x)
    y_ref, _ = orc.cusparse_spmv(A, x, alg=alg)
    assert float((y - y_ref).abs().max()) <= 1e-12 * float(y_ref.abs().max())


@pytest.mark.parametrize("name", sorted(SPMV_MATS))
def test_spmv_fp32_matches_cusparse(handle, name):
    a = SPMV_MATS[name]()
    A = DeviceCSR(a, "s")
    x64 = orc.dlarnv(a.n) - 0.5
    x = torch.from_numpy(x64.astype(np.float32)).cuda()
    y = handle.spmv(A, x).cpu().numpy().astype(np.float64)
    y_ref = orc.cusparse_spmv(A, x)[0].cpu().numpy().astype(np.float64)
    sums, lens = _row_abs_sums(a, x64)
    assert np.all(np.abs(y - y_ref) <= 2 * (lens + 2) * 2.0 ** -24 * sums + 1e-30)


def test_spmv_mixed_between_fp32_and_fp64_cusparse(handle):
    """mixed (fp32 A, fp64 x / accumulate) has no cuSPARSE counterpart: it must equal the fp64 cuSPARSE product of
    the fp32-rounded matrix to fp64 rounding."""
    a = host.gen_stencil27(32, seed=2)
    a32 = host.CSR(a.m, a.n, a.nnz, a.rowptr, a.colidx, a.values.astype(np.float32).astype(np.float64), a.base)
    x = torch.from_numpy(orc.dlarnv(a.n) - 0.5).cuda()
    y = handle.spmv(DeviceCSR(a, "m"), x)
    y_ref, _ = orc.cusparse_spmv(DeviceCSR(a32, "d"), x)
    assert float((y - y_ref).abs().max()) <= 1e-12 * float(y_ref.abs().max())


ILU_MATS = {
    "lap2d_60": lambda: host.gen_laplace2d(60),
    "femlike_20": lambda: host.gen_femlike(20),
    "stencil27_14": lambda: host.gen_stencil27(14, seed=2),
    "powerlaw_5k": lambda: host.gen_powerlaw(5000, 2.2, 48, 2, 900, 4),
}


def _lu_bound(a, lu):
    import scipy.sparse as sp
    rp, ci = a.rowptr.astype(np.int64), a.colidx.astype(np.int64)
    rows = np.repeat(np.arange(a.m), np.diff(rp))
    low = ci < rows
    lu = lu.astype(np.float64)
    L = sp.csr_matrix((np.where(low, np.abs(lu), 0.0), ci, rp), shape=(a.m, a.m)) + sp.identity(a.m, format="csr")
    U = sp.csr_matrix((np.where(low, 0.0, np.abs(lu)), ci, rp), shape=(a.m, a.m))
    return np.asarray((L @ U).tocsr()[rows, ci]).ravel(), int(np.diff(rp).max())


@pytest.mark.parametrize("name", sorted(ILU_MATS))
@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
@pytest.mark.parametrize("level", [0, 1])
def test_ilu0_and_solves_match_cusparse(handle, name, dtype, level):
    a = ILU_MATS[name]()
    prec = "d" if dtype == torch.float64 else "s"
    eps = 2.0 ** -53 if prec == "d" else 2.0 ** -24
    A = DeviceCSR(a, prec)
    ref_vals = A.values.clone()
    s_piv, n_piv, _, _ = orc.cusparse_ilu0(A, ref_vals, level=level)
    st, piv = handle.ilu0_factor(A)
    assert st == 0 and piv == -1 and s_piv == -1 and n_piv == -1
    lu, lu_ref = A.values.cpu().numpy(), ref_vals.cpu().numpy()
    bound, longest = _lu_bound(a, lu_ref)
    assert np.all(np.abs(lu.astype(np.float64) - lu_ref.astype(np.float64)) <= 4 * (longest + 2) * eps * bound)
    # both solves on cuSPARSE's factors, so only the solves differ
    b = torch.from_numpy(orc.dlarnv(a.m) + 0.5).cuda().to(dtype)
    z, y = handle.sptrsv_lu(A, ref_vals, b)
    z_ref, y_ref, _, _ = orc.cusparse_spsv_lu(A, ref_vals, b)
    tol = 1e-10 if prec == "d" else 2e-4
    assert float((z - z_ref).abs().max()) <= tol * float(z_ref.abs().max())
    assert float((y - y_ref).abs().max()) <= tol * float(y_ref.abs().max())


def test_zero_pivot_reports_match_cusparse(handle):
    # structural: b1_ss has no A(0,0) -- the harness prints "A(0,0) is missing" (ilu0.cu:221-226)
    a = host.load_matrix_market(os.path.join(GOLD, "b1_ss.mtx"))
    A = DeviceCSR(a, "d")
    s_piv, _, _, _ = orc.cusparse_ilu0(A, A.values.clone())
    st, row = handle.ilu0_analysis(A)
    assert row == s_piv == 0
    # numerical: [[1,1],[1,1]] -> U(1,1) == 0 -- "L(1,1) is zero" (ilu0.cu:278-282)
    a = host.CSR(2, 2, 4, np.array([0, 2, 4], np.int32), np.array([0, 1, 0, 1], np.int32), np.array([1.0, 1.0, 1.0, 1.0]))
    A = DeviceCSR(a, "d")
    ref_vals = A.values.clone()
    _, n_piv, _, _ = orc.cusparse_ilu0(A, ref_vals)
    st, row = handle.ilu0_factor(A)
    assert row == n_piv == 1
    assert torch.equal(A.values, ref_vals)


def test_config3_full_size_matches_cusparse(handle):
    a = host.gen_femlike(47, 0.58, 3)
    A = DeviceCSR(a, "d")
    ref_vals = A.values.clone()
    orc.cusparse_ilu0(A, ref_vals, level=1)
    handle.ilu0_factor(A)
    lu, lu_ref = A.values.cpu().numpy(), ref_vals.cpu().numpy()
    bound, longest = _lu_bound(a, lu_ref)
    assert np.all(np.abs(lu - lu_ref) <= 4 * (longest + 2) * 2.0 ** -53 * bound)
    b = torch.ones(a.m, dtype=torch.float64, device="cuda")
    z, y = handle.sptrsv_lu(A, ref_vals, b)
    z_ref, y_ref, _, _ = orc.cusparse_spsv_lu(A, ref_vals, b)
    assert float((y - y_ref).abs().max()) <= 1e-10 * float(y_ref.abs().max())


def test_config2_full_size_matches_cusparse(handle):
    a = host.gen_stencil27(100, seed=2)
    x64 = orc.dlarnv(a.n) - 0.5
    for prec, tol in (("d", 1e-12), ("s", 2e-5)):
        A = DeviceCSR(a, prec)
        x = torch.from_numpy(x64).cuda().to(A.values.dtype)
        y = handle.spmv(A, x)
        y_ref, _ = orc.cusparse_spmv(A, x)
        assert float((y - y_ref).abs().max()) <= tol * float(y_ref.abs().max())


@pytest.mark.parametrize("name", ["stencil27_40", "femlike_24", "powerlaw_20k"])
def test_transposed_product_matches_cusparse(handle, name):
    """spmv_csr_t_d/s against cusparseSpMV(CUSPARSE_OPERATION_TRANSPOSE) on the same arrays"""
    a = SPMV_MATS[name]()
    x64 = orc.dlarnv(a.m) - 0.5
    for prec, tol in (("d", 1e-12), ("s", 3e-5)):
        A = DeviceCSR(a, prec)
        x = torch.from_numpy(x64).cuda().to(A.values.dtype)
        y = handle.spmv_t(A, x)
        y_ref, _ = orc.cusparse_spmv(A, x, transpose=True)
        assert float((y - y_ref).abs().max()) <= tol * float(y_ref.abs().max())
