"""CSC / transpose path.  The reference gets a CSC by loading the file with transpose = 1
(loadMatrixMarket.cpp:79-81,124): the device transpose must reproduce THAT output bit for bit (indices and values),
and y = A^T x must equal the oracle's product with the loader-transposed matrix."""
import os

import numpy as np
import pytest
import torch

import orc
from respasol_b200 import host
from respasol_b200.gpu import DeviceCSR

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")


def via_file(tmp_path, a, name):
    """(A, A^T) as the loader returns them from one Matrix Market file"""
    p = tmp_path / (name + ".mtx")
    host.write_matrix_market(p, a)
    return host.load_matrix_market(p, a.base, 0), host.load_matrix_market(p, a.base, 1)


MATS = {
    "femlike_14": lambda: host.gen_femlike(14),
    "stencil27_slab": lambda: host.gen_stencil27(12, 12, 30, 10, 20, seed=2),  # rectangular: 1440 x 4320
    "powerlaw_30k": lambda: host.gen_powerlaw(30000, 2.2, 64, 3, 9000, 4),
    "lap2d_90": lambda: host.gen_laplace2d(90),
}


@pytest.mark.parametrize("name", sorted(MATS))
@pytest.mark.parametrize("base", [0, 1])
def test_device_transpose_equals_loader_transpose(handle, tmp_path, name, base):
    g = MATS[name]()
    g = host.CSR(g.m, g.n, g.nnz, g.rowptr + base, g.colidx + base, g.values, base)
    a, at = via_file(tmp_path, g, name)
    for prec, vt in (("d", np.float64), ("s", np.float32)):
        A = DeviceCSR(a, prec)
        colptr, rowidx, vals = handle.csr2csc(A)
        assert np.array_equal(colptr.cpu().numpy(), at.rowptr)
        assert np.array_equal(rowidx.cpu().numpy(), at.colidx[:at.stored])
        assert np.array_equal(vals.cpu().numpy(), at.values[:at.stored].astype(vt))


def test_reference_fixtures_transpose(handle):
    for f in ("b1_ss.mtx", "rect.mtx", "empty_rows.mtx", "one.mtx"):
        a = host.load_matrix_market(os.path.join(GOLD, f), 0, 0)
        at = host.load_matrix_market(os.path.join(GOLD, f), 0, 1)
        colptr, rowidx, vals = handle.csr2csc(DeviceCSR(a, "d"))
        assert np.array_equal(colptr.cpu().numpy(), at.rowptr), f
        assert np.array_equal(rowidx.cpu().numpy(), at.colidx[:at.stored]), f
        assert np.array_equal(vals.cpu().numpy(), at.values[:at.stored]), f


@pytest.mark.parametrize("name", sorted(MATS))
def test_transposed_product_matches_oracle(handle, tmp_path, name):
    a, at = via_file(tmp_path, MATS[name](), name)
    x = orc.dlarnv(a.m) - 0.5
    y0 = orc.dlarnv(a.n) - 0.5
    for prec, xt, tol in (("d", torch.float64, 1e-12), ("s", torch.float32, 3e-5)):
        A = DeviceCSR(a, prec)
        xd = torch.from_numpy(x).cuda().to(xt)
        ref = orc.spmv(at, x, prec).astype(np.float64)
        y = handle.spmv_t(A, xd).cpu().numpy().astype(np.float64)
        assert np.abs(y - ref).max() <= tol * np.abs(ref).max()
        # alpha / beta, and a second call after the values changed (values are re-gathered, the pattern is cached)
        A.values.mul_(2.0)
        yd = torch.from_numpy(y0).cuda().to(xt)
        y2 = handle.spmv_t(A, xd, yd, alpha=0.5, beta=-1.0).cpu().numpy().astype(np.float64)
        ref2 = ref - y0.astype(np.float32 if prec == "s" else np.float64).astype(np.float64)
        assert np.abs(y2 - ref2).max() <= 4 * tol * max(np.abs(ref2).max(), 1.0)
        # (A^T)^T x = A x
        z = torch.from_numpy(orc.dlarnv(a.n) - 0.5).cuda().to(xt)
        lhs = float((handle.spmv_t(A, xd) * z).sum())   # <A^T x, z>
        rhs = float((xd * handle.spmv(A, z)).sum())     # <x, A z>
        assert abs(lhs - rhs) <= (1e-10 if prec == "d" else 2e-3) * max(abs(rhs), 1.0)


def test_transpose_config2_full_size(handle):
    a = host.gen_stencil27(100, seed=2)
    A = DeviceCSR(a, "d")
    x = torch.from_numpy(orc.dlarnv(a.m) - 0.5).cuda()
    z = torch.from_numpy(orc.dlarnv(a.n) + 0.25).cuda()
    lhs = float((handle.spmv_t(A, x) * z).sum())
    rhs = float((x * handle.spmv(A, z)).sum())
    assert abs(lhs - rhs) <= 1e-10 * abs(rhs)
    colptr, rowidx, vals = handle.csr2csc(A)
    # transposing twice gives the matrix back, bit for bit
    At = object.__new__(DeviceCSR)
    At.uid, At.m, At.n, At.base, At.precision, At.stored = -1, a.n, a.m, 0, "d", a.stored
    At.rowptr, At.colidx, At.values, At.device = colptr, rowidx, vals, A.device
    DeviceCSR._next_uid += 1
    At.uid = DeviceCSR._next_uid
    rp2, ci2, v2 = handle.csr2csc(At)
    assert torch.equal(rp2, A.rowptr) and torch.equal(ci2, A.colidx) and torch.equal(v2, A.values)


def test_transposed_product_value_modes(handle):
    """RSG_OPT_TRANSPOSE_STATIC: 0 = values read through the permutation inside the product (one launch), 2 = gather +
    product (two launches), 1 = the caller promises static values: gathered once per (pattern, values pointer), then one
    launch.  All three agree to rounding, each is deterministic, and modes 0 / 2 see a change of the values."""
    a = host.gen_stencil27(24, seed=5)
    A = DeviceCSR(a, "d")
    x = torch.from_numpy(orc.dlarnv(a.m) - 0.5).cuda()
    scale = 1e-13 * float(handle.spmv_t(A, x).abs().max())

    def run(mode, M=A):
        handle.set_option("TRANSPOSE_STATIC", mode)
        try:
            l0 = handle.info("KERNEL_LAUNCHES")
            y = handle.spmv_t(M, x).clone()
            return y, handle.info("KERNEL_LAUNCHES") - l0
        finally:
            handle.set_option("TRANSPOSE_STATIC", 0)
    y0, n0 = run(0)
    y0b, _ = run(0)
    y2, n2 = run(2)
    y1, _ = run(1)
    y1b, n1 = run(1)
    assert n0 == 1 and n2 == 2 and n1 == 1
    assert torch.equal(y0, y0b) and torch.equal(y1, y1b) and torch.equal(y1, y2)   # 1 and 2 run the same product kernel
    assert float((y0 - y2).abs().max()) <= scale
    B = DeviceCSR(a, "d")                     # another values pointer on another pattern: gathered afresh under mode 1
    B.values.mul_(4.0)                        # a power of two: the scaled product is exact
    assert torch.equal(run(1, B)[0], 4.0 * y1)
    A.values.mul_(2.0)
    assert torch.equal(run(0)[0], 2.0 * y0) and torch.equal(run(2)[0], 2.0 * y2)
    # a pattern with a column too long for the fused kernel (the transpose of a matrix with two 5000-entry rows) falls back
    # to gather + product on its own
    import scipy.sparse as sp
    p = host.gen_powerlaw(6000, 2.2, 48, 2, 5000, 4)
    S = sp.csr_matrix((p.values, p.colidx.astype(np.int64) - p.base, p.rowptr.astype(np.int64) - p.base), shape=(p.m, p.n))
    St = S.T.tocsr()
    St.sort_indices()
    q = host.CSR(p.n, p.m, St.nnz, St.indptr.astype(np.int32), St.indices.astype(np.int32), St.data.copy(), 0)
    Q = DeviceCSR(q, "d")
    xq = torch.from_numpy(orc.dlarnv(q.m) - 0.5).cuda()
    handle.spmv_t(Q, xq)                      # builds the plans
    l0 = handle.info("KERNEL_LAUNCHES")
    yq = handle.spmv_t(Q, xq)
    assert handle.info("KERNEL_LAUNCHES") - l0 >= 2 and handle.info("SPMV_MAX_ROW") >= 5000
    ref = orc.spmv(p, xq.cpu().numpy(), "d")  # (S^T)^T x = S x
    assert np.abs(yq.cpu().numpy() - ref).max() <= 1e-12 * np.abs(ref).max()


def test_transpose_edge_cases(handle):
    # no stored entries: A^T x = 0 (or beta*y), csr2csc gives an all-base colptr
    e = host.CSR(3, 5, 0, np.zeros(4, np.int32), np.zeros(0, np.int32), np.zeros(0))
    E = DeviceCSR(e, "d")
    x = torch.ones(3, dtype=torch.float64, device="cuda")
    y = torch.full((5,), 7.0, dtype=torch.float64, device="cuda")
    out = handle.spmv_t(E, x, y, alpha=1.0, beta=0.5)
    assert torch.equal(out.cpu(), torch.full((5,), 3.5, dtype=torch.float64))
    colptr, rowidx, vals = handle.csr2csc(E)
    assert colptr.cpu().tolist() == [0] * 6 and rowidx.numel() == 0 and vals.numel() == 0
    # one-based, rectangular, a single row
    r = host.CSR(1, 4, 3, np.array([1, 4], np.int32), np.array([1, 3, 4], np.int32), np.array([2.0, -1.0, 4.0]), 1)
    R = DeviceCSR(r, "d")
    out = handle.spmv_t(R, torch.tensor([3.0], dtype=torch.float64, device="cuda"))
    assert out.cpu().tolist() == [6.0, 0.0, -3.0, 12.0]
    colptr, rowidx, vals = handle.csr2csc(R)
    assert colptr.cpu().tolist() == [1, 2, 2, 3, 4] and rowidx.cpu().tolist() == [1, 1, 1] and vals.cpu().tolist() == [2.0, -1.0, 4.0]
