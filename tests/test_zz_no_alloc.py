"""spmv_csr_d_noalloc -- the allocation-free fp64 SpMV (plan in a file-scope device array, no cudaMalloc) -- against the
existing entry point spmv_csr_d, through the C ABI, on two calls in a row (the second on another matrix: the array keeps the
first call's plan and must be rebuilt).  No counterpart in the reference; see include/reaspasol_gpu.h.
Named test_zz_* so that pytest collects it last."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

TOL = 1e-12  # fp64 SpMV tolerance of tests/test_gpu_spmv.py: |y - y_ref|_inf <= 1e-12 * |y_ref|_inf
SIZE_LIMIT = 10


def _setup():
    import torch
    from respasol_b200 import _lib, host
    from respasol_b200.gpu import DeviceCSR, Handle
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    return torch, _lib.gpu_lib(), host, DeviceCSR, Handle, dev


def _p(t):
    return C.c_void_p(t.data_ptr() if t is not None and t.numel() else None)


def _noalloc(lib, A, x, y, alpha=1.0, beta=0.0, stream=None):
    return lib.spmv_csr_d_noalloc(A.m, A.n, A.stored, A.base, _p(A.rowptr), _p(A.colidx), _p(A.values), alpha, _p(x), beta, _p(y),
                                  C.c_void_p(stream))


def test_noalloc_matches_spmv_csr_d_two_calls_in_a_row():
    torch, lib, host, DeviceCSR, Handle, dev = _setup()
    rng = np.random.default_rng(7)
    mats = [host.gen_stencil27(24, seed=2),                       # 13 824 rows, 27-point rows (two lanes per row)
            host.gen_laplace2d(150),                              # 22 500 short rows: another tile size, another plan
            host.gen_powerlaw(6000, 2.2, 16, 2, 5000, 4)]         # ragged rows incl. rows longer than a stage (CTA-per-row path)
    h = Handle(0)
    try:
        for rep in range(2):       # the whole sequence twice: every call finds the previous call's plan in the array
            for a in mats:
                A = DeviceCSR(a, "d", dev)
                x = torch.from_numpy(rng.standard_normal(a.n)).to(dev)
                y0 = torch.from_numpy(rng.standard_normal(a.m)).to(dev)
                for alpha, beta in ((1.0, 0.0), (-0.75, 0.5)):
                    ref = y0.clone()
                    h.spmv(A, x, ref, alpha=alpha, beta=beta)
                    got = y0.clone()
                    assert _noalloc(lib, A, x, got, alpha, beta) == 0              # (the very first call also loads the kernels)
                    torch.cuda.synchronize()
                    free0 = torch.cuda.mem_get_info(dev)[0]
                    assert _noalloc(lib, A, x, got.copy_(y0), alpha, beta) == 0
                    assert _noalloc(lib, A, x, got.copy_(y0), alpha, beta) == 0  # again, same stream, back to back
                    torch.cuda.synchronize()
                    assert torch.cuda.mem_get_info(dev)[0] >= free0 - (2 << 20)  # nothing allocated (allow the driver's own noise)
                    scale = float(ref.abs().max())
                    assert float((got - ref).abs().max()) <= TOL * max(scale, 1e-300), (rep, a.m, alpha, beta)
    finally:
        h.close()


def test_noalloc_on_a_side_stream_and_base_one():
    torch, lib, host, DeviceCSR, Handle, dev = _setup()
    a = host.gen_stencil27(16, seed=3)
    a1 = host.CSR(a.m, a.n, a.nnz, a.rowptr + 1, a.colidx + 1, a.values, 1)
    h = Handle(0)
    try:
        A = DeviceCSR(a1, "d", dev)
        x = torch.from_numpy(np.random.default_rng(1).standard_normal(a.n)).to(dev)
        ref = h.spmv(A, x)
        torch.cuda.synchronize()
        st = torch.cuda.Stream(dev)
        got = torch.full((a.m,), float("nan"), dtype=torch.float64, device=dev)
        torch.cuda.synchronize()
        assert _noalloc(lib, A, x, got, stream=st.cuda_stream) == 0
        st.synchronize()
        assert float((got - ref).abs().max()) <= TOL * float(ref.abs().max())
    finally:
        h.close()


def test_noalloc_rejects_sizes_beyond_the_scratch_and_bad_arguments():
    torch, lib, host, DeviceCSR, Handle, dev = _setup()
    one = torch.zeros(8, dtype=torch.int32, device=dev)
    v = torch.zeros(8, dtype=torch.float64, device=dev)
    # rows beyond RSG_NOALLOC_MAX_ROWS: refused before anything is launched or read
    rc = lib.spmv_csr_d_noalloc((1 << 20) + 1, 8, 8, 0, _p(one), _p(one), _p(v), 1.0, _p(v), 0.0, _p(v), C.c_void_p(None))
    assert rc == SIZE_LIMIT and b"scratch" in lib.rsg_status_string(rc)
    # tiles beyond RSG_NOALLOC_MAX_TILES (4*m + nnz > 65536 * 2048)
    rc = lib.spmv_csr_d_noalloc(1 << 20, 8, 200_000_000, 0, _p(one), _p(one), _p(v), 1.0, _p(v), 0.0, _p(v), C.c_void_p(None))
    assert rc == SIZE_LIMIT
    assert lib.spmv_csr_d_noalloc(4, 8, 8, 2, _p(one), _p(one), _p(v), 1.0, _p(v), 0.0, _p(v), C.c_void_p(None)) == 1   # base
    assert lib.spmv_csr_d_noalloc(4, 8, 8, 0, None, _p(one), _p(v), 1.0, _p(v), 0.0, _p(v), C.c_void_p(None)) == 1      # rowptr NULL
    mis = C.c_void_p(one.data_ptr() + 4)
    assert lib.spmv_csr_d_noalloc(4, 8, 4, 0, _p(one), mis, _p(v), 1.0, _p(v), 0.0, _p(v), C.c_void_p(None)) == 1       # unaligned colidx
    assert lib.spmv_csr_d_noalloc(0, 8, 0, 0, _p(one), None, None, 1.0, None, 0.0, None, C.c_void_p(None)) == 0         # empty matrix
    torch.cuda.synchronize()
