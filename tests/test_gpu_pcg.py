"""rsg_pcg_d (the Krylov loop around the hot path) against the oracle's PCG on the same inputs.

Dot products are summed in a different order on the GPU, so the iterates agree to rounding only: the iteration
count must match the oracle's within one, the solutions to 1e-6 relative (rtol is 1e-9), and the TRUE residual
|b - A x| / |b| must meet 10 rtol.  The preconditioned runs must take fewer iterations than the plain one, and the
fp32-factor run (mixed precision, the reference README's theme) about as many as the fp64-factor run."""
import ctypes as C

import numpy as np
import pytest
import torch

import orc
from respasol_b200 import host
from respasol_b200.gpu import DeviceCSR

pytestmark = pytest.mark.gpu

def sym_femlike(N, p=0.58, seed=3):
    """gen_femlike is structurally symmetric with unsymmetric values: its symmetric part is SPD (diagonally dominant)"""
    import scipy.sparse as sp
    g = host.gen_femlike(N, p, seed)
    G = sp.csr_matrix((g.values, g.colidx, g.rowptr), shape=(g.m, g.n))
    M = ((G + G.T) * 0.5).tocsr()
    M.sort_indices()
    return host.CSR(g.m, g.n, M.nnz, M.indptr.astype(np.int32), M.indices.astype(np.int32), M.data.copy(), 0)


MATS = {
    "lap2d_80": lambda: host.gen_laplace2d(80),
    "sym_femlike_22": lambda: sym_femlike(22),
    "stencil27_24": lambda: host.gen_stencil27(24, seed=0),
}


@pytest.mark.parametrize("name", sorted(MATS))
def test_pcg_matches_oracle(handle, name):
    a = MATS[name]()
    A = DeviceCSR(a, "d")
    xs = orc.dlarnv(a.n) - 0.5
    b = orc.spmv(a, xs, "d")
    bd = torch.from_numpy(b).cuda()
    rtol = 1e-9
    iters = {}
    for pre in (None, "d", "s"):
        lu = lu_ref = None
        if pre:
            F = DeviceCSR(a, pre)
            st, _ = handle.ilu0_factor(F)
            assert st == 0
            lu = F.values
            lu_ref = orc.ilu0(a, np.float64 if pre == "d" else np.float32)[2]
        x, it, rel = handle.pcg(A, bd, lu=lu, max_iters=2000, rtol=rtol)
        x_ref, it_ref, rel_ref = orc.pcg(a, b, lu_ref, max_iters=2000, rtol=rtol)
        assert abs(it - it_ref) <= 1 and it < 2000, (pre, it, it_ref)
        assert rel <= rtol
        xh = x.cpu().numpy()
        assert np.abs(xh - x_ref).max() <= 1e-6 * np.abs(x_ref).max()
        true_res = np.linalg.norm(b - orc.spmv(a, xh, "d")) / np.linalg.norm(b)
        assert true_res <= 10 * rtol
        assert np.abs(xh - xs).max() <= 1e-5 * np.abs(xs).max()
        iters[pre] = it
    assert iters["d"] < iters[None] and iters["s"] <= iters["d"] + 2


def test_pcg_initial_guess_and_limits(handle):
    a = host.gen_laplace2d(40)
    A = DeviceCSR(a, "d")
    xs = orc.dlarnv(a.n)
    b = torch.from_numpy(orc.spmv(a, xs, "d")).cuda()
    # exact initial guess: zero iterations
    x, it, rel = handle.pcg(A, b, x=torch.from_numpy(xs).cuda(), rtol=1e-10)
    assert it == 0 and rel <= 1e-10
    # iteration cap is honoured and reported
    x, it, rel = handle.pcg(A, b, max_iters=3, rtol=1e-14)
    _, it_ref, rel_ref = orc.pcg(a, b.cpu().numpy(), None, max_iters=3, rtol=1e-14)
    assert it == it_ref == 3 and abs(rel - rel_ref) <= 1e-9 * rel_ref


def test_pcg_config3_full_size(handle):
    """symmetric part of config 3 (FEM-like, 103 823 rows): fp32 ILU(0) inside fp64 CG reaches 1e-10 like the fp64 factors do."""
    a = sym_femlike(47, 0.58, 3)
    A = DeviceCSR(a, "d")
    xs = orc.dlarnv(a.n) - 0.5
    b = torch.from_numpy(orc.spmv(a, xs, "d")).cuda()
    its = {}
    for pre in ("d", "s"):
        F = DeviceCSR(a, pre)
        handle.ilu0_factor(F)
        x, it, rel = handle.pcg(A, b, lu=F.values, max_iters=500, rtol=1e-10)
        assert rel <= 1e-10 and it < 500
        assert float((x - torch.from_numpy(xs).cuda()).abs().max()) <= 1e-7
        its[pre] = it
    assert abs(its["d"] - its["s"]) <= 2


NONSYM = {
    "femlike_22": lambda: host.gen_femlike(22),                 # symmetric pattern, unsymmetric values
    "stencil27_24_perturbed": lambda: host.gen_stencil27(24, seed=2),
    "lap2d_80": lambda: host.gen_laplace2d(80),                 # SPD works too
}


@pytest.mark.parametrize("name", sorted(NONSYM))
def test_bicgstab_matches_oracle(handle, name):
    """BiCGStab is less regular than CG: iteration counts of two correctly rounded runs may differ by a few; both
    must converge, agree on x to 1e-6 and meet the true-residual bound."""
    a = NONSYM[name]()
    A = DeviceCSR(a, "d")
    xs = orc.dlarnv(a.n) - 0.5
    b = orc.spmv(a, xs, "d")
    bd = torch.from_numpy(b).cuda()
    rtol = 1e-9
    iters = {}
    for pre in (None, "d", "s"):
        lu = lu_ref = None
        if pre:
            F = DeviceCSR(a, pre)
            st, _ = handle.ilu0_factor(F)
            assert st == 0
            lu = F.values
            lu_ref = orc.ilu0(a, np.float64 if pre == "d" else np.float32)[2]
        x, it, rel = handle.bicgstab(A, bd, lu=lu, max_iters=2000, rtol=rtol)
        x_ref, it_ref, rel_ref = orc.bicgstab(a, b, lu_ref, max_iters=2000, rtol=rtol)
        assert it < 2000 and it_ref < 2000 and abs(it - it_ref) <= max(3, it_ref // 10), (pre, it, it_ref)
        assert rel <= rtol
        xh = x.cpu().numpy()
        assert np.abs(xh - x_ref).max() <= 1e-6 * np.abs(x_ref).max()
        assert np.linalg.norm(b - orc.spmv(a, xh, "d")) / np.linalg.norm(b) <= 10 * rtol
        iters[pre] = it
    assert iters["d"] < iters[None] and iters["s"] <= iters["d"] + 3


def test_bicgstab_config3_full_size(handle):
    """config 3 as generated (FEM-like, 103 823 rows, unsymmetric values): ILU(0)-preconditioned BiCGStab, fp64 and
    fp32 factors, reaches 1e-10."""
    a = host.gen_femlike(47, 0.58, 3)
    A = DeviceCSR(a, "d")
    xs = orc.dlarnv(a.n) - 0.5
    b = torch.from_numpy(orc.spmv(a, xs, "d")).cuda()
    its = {}
    for pre in ("d", "s"):
        F = DeviceCSR(a, pre)
        handle.ilu0_factor(F)
        x, it, rel = handle.bicgstab(A, b, lu=F.values, max_iters=500, rtol=1e-10)
        assert rel <= 1e-10 and it < 500
        assert float((x - torch.from_numpy(xs).cuda()).abs().max()) <= 1e-7
        its[pre] = it
    assert abs(its["d"] - its["s"]) <= 3


@pytest.mark.parametrize("name", sorted(NONSYM))
def test_gmres_matches_oracle(handle, name):
    """Flexible GMRES(30): same Arnoldi steps as the oracle (dot products summed in a different, fixed order), so the
    iteration counts agree to +-2, the solutions to 1e-6 and the true residual meets the bound; several restarts without
    a preconditioner."""
    a = NONSYM[name]()
    A = DeviceCSR(a, "d")
    xs = orc.dlarnv(a.n) - 0.5
    b = orc.spmv(a, xs, "d")
    bd = torch.from_numpy(b).cuda()
    rtol = 1e-9
    iters = {}
    for pre in (None, "d", "s"):
        lu = lu_ref = None
        if pre:
            F = DeviceCSR(a, pre)
            st, _ = handle.ilu0_factor(F)
            assert st == 0
            lu = F.values
            lu_ref = orc.ilu0(a, np.float64 if pre == "d" else np.float32)[2]
        x, it, rel = handle.gmres(A, bd, lu=lu, max_iters=2000, rtol=rtol)
        x_ref, it_ref, rel_ref = orc.gmres(a, b, lu_ref, max_iters=2000, rtol=rtol)
        assert it < 2000 and it_ref < 2000 and abs(it - it_ref) <= 2, (pre, it, it_ref)
        assert rel <= rtol
        xh = x.cpu().numpy()
        assert np.abs(xh - x_ref).max() <= 1e-6 * np.abs(x_ref).max()
        assert np.linalg.norm(b - orc.spmv(a, xh, "d")) / np.linalg.norm(b) <= 10 * rtol
        iters[pre] = it
    assert iters["d"] < iters[None] and iters["s"] <= iters["d"] + 2
    # iteration cap inside a cycle: x carries the partial cycle, the estimate is the true residual
    x7, it7, rel7 = handle.gmres(A, bd, max_iters=7, rtol=rtol)
    assert it7 == 7 and 0 < rel7 < 1
    assert abs(np.linalg.norm(b - orc.spmv(a, x7.cpu().numpy(), "d")) / np.linalg.norm(b) - rel7) <= 1e-9


def test_gmres_config3_full_size(handle):
    """config 3 as generated: ILU(0)-preconditioned flexible GMRES(30), fp64 and fp32 factors, reaches 1e-10."""
    a = host.gen_femlike(47, 0.58, 3)
    A = DeviceCSR(a, "d")
    xs = orc.dlarnv(a.n) - 0.5
    b = torch.from_numpy(orc.spmv(a, xs, "d")).cuda()
    its = {}
    for pre in ("d", "s"):
        F = DeviceCSR(a, pre)
        handle.ilu0_factor(F)
        x, it, rel = handle.gmres(A, b, lu=F.values, max_iters=500, rtol=1e-10)
        assert rel <= 1e-10 and it < 500
        assert float((x - torch.from_numpy(xs).cuda()).abs().max()) <= 1e-7
        its[pre] = it
    assert abs(its["d"] - its["s"]) <= 2


@pytest.mark.parametrize("solver", ["pcg", "bicgstab", "gmres"])
def test_krylov_edge_cases(handle, solver):
    from respasol_b200.gpu import RsgError
    run = getattr(handle, solver)
    # empty system: nothing to do, no device work
    e = host.CSR(0, 0, 0, np.zeros(1, np.int32), np.zeros(0, np.int32), np.zeros(0))
    E = DeviceCSR(e, "d")
    l0 = handle.info("KERNEL_LAUNCHES")
    x, it, rel = run(E, torch.zeros(0, dtype=torch.float64, device="cuda"))
    assert it == 0 and rel == 0.0 and handle.info("KERNEL_LAUNCHES") == l0
    # zero right-hand side with a zero initial guess: converged before the first iteration
    a = sym_femlike(8) if solver == "pcg" else host.gen_femlike(8)
    A = DeviceCSR(a, "d")
    x, it, rel = run(A, torch.zeros(a.m, dtype=torch.float64, device="cuda"))
    assert it == 0 and rel == 0.0 and float(x.abs().max()) == 0.0
    # max_iters = 0: the initial residual is reported, x untouched
    b = torch.from_numpy(orc.dlarnv(a.m)).cuda()
    x, it, rel = run(A, b, max_iters=0)
    assert it == 0 and abs(rel - 1.0) <= 1e-14 and float(x.abs().max()) == 0.0
    # a precision letter the interface does not know
    it_c, rel_c = C.c_int(0), C.c_double(0.0)
    fn = getattr(handle.lib, "rsg_%s_d" % solver)
    st = fn(handle.h, A.m, A.stored, A.base, handle._p(A.rowptr), handle._p(A.colidx), handle._p(A.values), handle._p(A.values),
            ord("q"), handle._p(b), handle._p(x), 5, 1e-8, C.byref(it_c), C.byref(rel_c))
    assert st == 1  # RSG_BAD_ARG
    with pytest.raises(RsgError):
        run(A, b, max_iters=-1)
