"""Multi-GPU row-block SpMV through the C ABI (rsg_spmv_multi_*: one process, N devices).

The N > 1 cases need N real devices (kernels that wait on each other must never share one GPU): they skip on a
one-GPU box and run under `gpurun --gpus 2` / on the 8-GPU box.  The one-rank case runs everywhere and pins the host
logic of the entry points (partition, relabelling, scatter / gather).  Full y against the oracle, bit-exact where the
rows are summed left to right like the oracle (window kernel, rows up to 40 entries)."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest
import torch

import orc
from respasol_b200 import host
from respasol_b200._lib import gpu_lib

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "respasol_b200", "bin")


def ndev():
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


def multi_product(a, x, ngpu, precision="d", row_starts=None, steps=3):
    lib = gpu_lib()
    M = C.c_void_p()
    rs = None if row_starts is None else np.ascontiguousarray(row_starts, dtype=np.int32)
    st = lib.rsg_spmv_multi_create(ngpu, None, precision.encode(), a.m, a.n, a.stored, a.base, a.rowptr.ctypes.data,
                                   a.colidx.ctypes.data, np.ascontiguousarray(a.values, dtype=np.float64).ctypes.data,
                                   None if rs is None else rs.ctypes.data, C.byref(M))
    assert st == 0, lib.rsg_status_string(st)
    xt = np.float32 if precision == "s" else np.float64
    xh = np.ascontiguousarray(x, dtype=xt)
    y = np.zeros(a.m, dtype=xt)
    try:
        assert lib.rsg_spmv_multi_set_x(M, xh.ctypes.data) == 0
        assert lib.rsg_spmv_multi_step(M, steps) == 0          # several steps: both x images, epochs 1..steps
        assert lib.rsg_spmv_multi_get_y(M, y.ctypes.data) == 0
        us = C.c_double(0)
        assert lib.rsg_spmv_multi_time(M, 5, C.byref(us)) == 0 and us.value > 0
        y2 = np.zeros(a.m, dtype=xt)
        assert lib.rsg_spmv_multi_get_y(M, y2.ctypes.data) == 0
        assert np.array_equal(y, y2)
        v = C.c_longlong(0)
        cuts = np.zeros(ngpu + 1, dtype=np.int32)
        assert lib.rsg_spmv_multi_info(M, 0, C.byref(v), cuts.ctypes.data) == 0 and v.value == ngpu
        assert cuts[0] == 0 and cuts[-1] == a.m and np.all(np.diff(cuts) > 0)
        if row_starts is None:
            assert np.array_equal(cuts, orc.row_partition(a, ngpu))  # the oracle's nnz-balanced cuts, bit-exact
    finally:
        assert lib.rsg_spmv_multi_destroy(M) == 0
    return y


@pytest.mark.parametrize("ngpu", [1, 2, 4, 8])
@pytest.mark.parametrize("precision", ["d", "s", "m"])
def test_multi_gpu_product_matches_the_oracle(ngpu, precision):
    if ndev() < ngpu:
        pytest.skip("needs %d devices" % ngpu)
    for a in (host.gen_stencil27(30, 30, 48, seed=2), host.gen_laplace2d(300), host.gen_femlike(20, 0.6, 5)):
        x = orc.dlarnv(a.n) - 0.5
        y = multi_product(a, x, ngpu, precision)
        ref = orc.spmv(a, x, precision)
        if precision == "d":
            assert np.array_equal(y, ref)
        else:
            assert np.abs(y.astype(np.float64) - ref.astype(np.float64)).max() <= (5e-5 if precision == "s" else 1e-12) * np.abs(ref).max()


@pytest.mark.parametrize("proto", ["pull", "packets", "flags"])
@pytest.mark.parametrize("ngpu", [2, 4])
def test_multi_gpu_one_directional_pattern_and_changing_x(ngpu, proto, monkeypatch):
    """An upwind (lower-bidiagonal-by-planes) pattern: rank r reads only from rank r-1, so rank 0 feeds a neighbour but
    reads nothing -- it must still not run ahead (its last tile awaits the flags).  x changes between steps, so a stale
    or early halo would show."""
    if ndev() < ngpu:
        pytest.skip("needs %d devices" % ngpu)
    n, band = 40000, 700
    rows = np.arange(n)
    rp = np.zeros(n + 1, dtype=np.int32)
    lens = np.where(rows < band, 1, 2)
    rp[1:] = np.cumsum(lens)
    ci = np.concatenate([[r] if r < band else [r - band, r] for r in range(n)]).astype(np.int32)
    a = host.CSR(n, n, int(rp[-1]), rp, ci, np.random.default_rng(1).uniform(0.5, 1.5, int(rp[-1])), 0)
    monkeypatch.setenv("RSG_DIST_EXCHANGE", proto)
    lib = gpu_lib()
    M = C.c_void_p()
    assert lib.rsg_spmv_multi_create(ngpu, None, b"d", n, n, a.stored, 0, rp.ctypes.data, ci.ctypes.data, a.values.ctypes.data, None, C.byref(M)) == 0
    try:
        v = C.c_longlong(-1)
        # packets need a mutual pattern (the arrival flags stay); pull and flags take any
        assert lib.rsg_spmv_multi_info(M, 4, C.byref(v), None) == 0 and v.value == {"pull": 2, "packets": 0, "flags": 0}[proto]  # (packets: not mutual)
        y = np.zeros(n)
        for it in range(6):
            x = orc.dlarnv(n) + it
            assert lib.rsg_spmv_multi_set_x(M, x.ctypes.data) == 0
            assert lib.rsg_spmv_multi_step(M, 1) == 0
            assert lib.rsg_spmv_multi_get_y(M, y.ctypes.data) == 0
            assert np.array_equal(y, orc.spmv(a, x, "d")), it
    finally:
        lib.rsg_spmv_multi_destroy(M)


@pytest.mark.parametrize("proto", ["pull", "packets", "packets-fused", "flags"])
@pytest.mark.parametrize("ngpu", [2, 4, 8])
def test_multi_gpu_exchange_protocols_changing_x(ngpu, proto, monkeypatch):
    """The three exchange protocols of rsg_spmv_multi_* on symmetric patterns: pull (the product's first CTAs fetch the halo --
    rsg_dist_enable_pull, the default), {value, step} packets unpacked by the halo tiles (rsg_dist_enable_packets) and push +
    arrival flags.  x changes every step and many steps run back to back, so a stale entry (the image of two steps ago), a
    torn packet or an early read would show; fp64 bit-exact."""
    if ndev() < ngpu:
        pytest.skip("needs %d devices" % ngpu)
    monkeypatch.setenv("RSG_DIST_EXCHANGE", proto)
    packets = {"flags": 0, "packets": 1, "packets-fused": 1, "pull": 2}[proto]
    lib = gpu_lib()
    for a, prec in ((host.gen_stencil27(30, 30, 48, seed=2), "d"), (host.gen_laplace2d(301), "d"), (host.gen_stencil27(20, 20, 40, seed=5), "s")):
        M = C.c_void_p()
        assert lib.rsg_spmv_multi_create(ngpu, None, prec.encode(), a.m, a.n, a.stored, a.base, a.rowptr.ctypes.data, a.colidx.ctypes.data,
                                         np.ascontiguousarray(a.values, dtype=np.float64).ctypes.data, None, C.byref(M)) == 0
        try:
            v = C.c_longlong(-1)
            assert lib.rsg_spmv_multi_info(M, 4, C.byref(v), None) == 0 and v.value == packets
            xt = np.float64 if prec == "d" else np.float32
            y = np.zeros(a.m, dtype=xt)
            for it in range(5):
                x = np.ascontiguousarray(orc.dlarnv(a.n) - 0.5 + 0.25 * it, dtype=xt)
                assert lib.rsg_spmv_multi_set_x(M, x.ctypes.data) == 0
                assert lib.rsg_spmv_multi_step(M, 1 + 2 * it) == 0   # odd and even step counts: both images, epochs far apart
                assert lib.rsg_spmv_multi_get_y(M, y.ctypes.data) == 0
                ref = orc.spmv(a, x.astype(np.float64), prec)
                if prec == "d":
                    assert np.array_equal(y, ref), (it, packets)
                else:
                    assert np.abs(y.astype(np.float64) - ref.astype(np.float64)).max() <= 5e-5 * np.abs(ref).max()
            us = C.c_double(0)
            assert lib.rsg_spmv_multi_time(M, 50, C.byref(us)) == 0 and us.value > 0
            assert lib.rsg_spmv_multi_get_y(M, y.ctypes.data) == 0
            if prec == "d":
                assert np.array_equal(y, ref)
        finally:
            assert lib.rsg_spmv_multi_destroy(M) == 0


@pytest.mark.parametrize("ngpu", [1, 2])
def test_harness_gpus_flag(tmp_path, ngpu):
    """bin/test_spmv --gpus N prints the reference's two stdout lines (GPU/spmv.cu:203-207,260)"""
    if ndev() < ngpu:
        pytest.skip("needs %d devices" % ngpu)
    p = tmp_path / "s27.mtx"
    host.write_matrix_market(p, host.gen_stencil27(24, seed=2))
    r = subprocess.run([os.path.join(BIN, "test_spmv"), str(p), "--gpus=%d" % ngpu, "--stats", "--repeat=20", "--warmup=2"],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    lines = r.stdout.strip().splitlines()
    assert re.fullmatch(r"DOUBLE PRECISION SPMV solve time \(microseconds\) = (\d+\.\d{6})", lines[0]), r.stdout
    m = re.fullmatch(r"Error= (\d\.\d{6}e[+-]\d{2})", lines[1])
    assert m and float(m.group(1)) <= 1e-12, r.stdout
    if ngpu > 1:
        assert "gpus=%d" % ngpu in lines[2]


@pytest.mark.parametrize("ngpu", [1, 2, 4, 8])
@pytest.mark.parametrize("precision", ["d", "m"])
def test_multi_gpu_chained_products(ngpu, precision):
    """x <- alpha * A * x, K times (rsg_spmv_multi_chain: boundary rows pushed by the tiles that computed them, three x images):
    the whole vector after K steps against the oracle's chain -- bit-exact in fp64, and again after continuing the chain."""
    if ndev() < ngpu:
        pytest.skip("needs %d devices" % ngpu)
    lib = gpu_lib()
    for a, alpha in ((host.gen_stencil27(30, 30, 48, seed=2), 1.0 / 52), (host.gen_laplace2d(300), 1.0 / 8), (host.gen_femlike(20, 0.6, 5), 0.25)):
        M = C.c_void_p()
        assert lib.rsg_spmv_multi_create(ngpu, None, precision.encode(), a.m, a.n, a.stored, a.base, a.rowptr.ctypes.data, a.colidx.ctypes.data,
                                         np.ascontiguousarray(a.values, dtype=np.float64).ctypes.data, None, C.byref(M)) == 0
        try:
            x0 = orc.dlarnv(a.n) - 0.5
            ref = x0.copy()
            got = np.zeros(a.n)
            us = C.c_double(0)
            done = 0
            for steps in (1, 2, 7):  # restart once, then continue the chain twice
                st = lib.rsg_spmv_multi_chain(M, x0.ctypes.data if done == 0 else None, steps, alpha, C.byref(us))
                assert st == 0, lib.rsg_status_string(st)
                for _ in range(steps):
                    ref = orc.spmv(a, ref, precision, alpha=alpha)
                done += steps
                assert lib.rsg_spmv_multi_get_x(M, got.ctypes.data) == 0
                if precision == "d":
                    assert np.array_equal(got, ref), (ngpu, done)
                else:
                    assert np.abs(got - ref).max() <= 1e-5 * np.abs(ref).max()
            # a fresh start must not see anything of the old chain
            st = lib.rsg_spmv_multi_chain(M, x0.ctypes.data, 3, alpha, None)
            assert st == 0
            assert lib.rsg_spmv_multi_get_x(M, got.ctypes.data) == 0
            r3 = x0.copy()
            for _ in range(3):
                r3 = orc.spmv(a, r3, precision, alpha=alpha)
            assert np.array_equal(got, r3) if precision == "d" else np.abs(got - r3).max() <= 1e-5 * np.abs(r3).max()
        finally:
            lib.rsg_spmv_multi_destroy(M)


# ---------------------------------------------------------------------------------------------------------------
# ONE device: a rank that exchanges with itself.  The block is a periodic 27-point stencil slab whose out-of-slab planes are
# halo columns fed from its own boundary planes, so every exchange protocol of rsg_dist_* (flags, packets on a side stream,
# packets pushed by the product, pull) runs through the same kernels and the same two-image / step-number logic as between
# GPUs -- on the single-GPU box the round's end tests on.
def _loopback_problem(nx=26, nz=20, seed=3):
    a = host.gen_stencil27(nx, nx, nz + 2, 1, nz + 1, seed=seed)  # rows of planes [1, nz+1), global columns of nz+2 planes
    plane = nx * nx
    m = a.m
    assert m == nz * plane
    hbase = (m + 31) // 32 * 32
    cols = a.colidx.astype(np.int64) - a.base
    new = cols - plane                                   # own planes -> local numbering
    below, above = cols < plane, cols >= (nz + 1) * plane
    # periodic wrap: plane 0 is fed from the slab's LAST plane, plane nz+1 from its FIRST one; send list = [first plane | last plane]
    new[above] = hbase + (cols[above] - (nz + 1) * plane)            # halo slot [0, plane): the first plane
    new[below] = hbase + plane + cols[below]                         # halo slot [plane, 2 plane): the last plane
    n_ext = hbase + 2 * plane
    blk = host.CSR(m, n_ext, a.stored, (a.rowptr - a.base).astype(np.int32), new.astype(np.int32), a.values, 0)
    send = np.concatenate([np.arange(plane), np.arange(m - plane, m)]).astype(np.int32)
    return blk, hbase, send, plane


def _loopback_reference(blk, hbase, send, x_local):
    x_ext = np.zeros(blk.n)
    x_ext[:blk.m] = x_local
    x_ext[hbase:hbase + len(send)] = x_local[send]
    return orc.spmv(blk, x_ext, "d")


@pytest.mark.parametrize("proto", ["flags", "packets", "packets-fused", "pull"])
def test_single_device_loopback_exchange(proto):
    from respasol_b200.gpu import DeviceCSR, Handle
    dev = torch.device("cuda", 0)
    blk, hbase, send, plane = _loopback_problem()
    h = Handle(0)
    try:
        A = DeviceCSR(blk, "d", dev)
        stride = (blk.n + 31) // 32 * 32
        ximg = torch.zeros(2 * stride, dtype=torch.float64, device=dev)
        flags = torch.zeros(128, dtype=torch.int32, device=dev)
        send_d = torch.from_numpy(send).to(dev)
        img = [ximg[p * stride:] for p in range(2)]
        slots = [[img[p].data_ptr() + 8 * hbase] for p in range(2)]          # this rank's own halo: "peer 0" is the rank itself
        pflags = [[flags[32 * p:].data_ptr()] for p in range(2)]
        d = h.dist_create(A, [t.data_ptr() for t in img], hbase, send_d, slots, pflags, [0], [len(send)],
                          [flags[32 * p:].data_ptr() for p in range(2)], 1)
        keep = []
        if proto.startswith("packets"):
            pk = torch.zeros(2 * 2 * plane * 2, dtype=torch.float64, device=dev)  # 2 images x 2 plane packets of 16 bytes
            keep.append(pk)
            mine = [pk.data_ptr() + 16 * 2 * plane * p for p in range(2)]
            assert h.dist_enable_packets(d, mine, [[mine[0]], [mine[1]]], side_stream=(proto == "packets"))
        elif proto == "pull":
            col = torch.from_numpy((hbase + np.arange(len(send))).astype(np.int32)).to(dev)
            keep += [col]
            assert h.dist_enable_pull(d, col, send_d, [[img[0].data_ptr()], [img[1].data_ptr()]], [flags[96:].data_ptr()], [0], [len(send)],
                                      flags[96:].data_ptr(), flags[97:].data_ptr())
        y = torch.zeros(blk.m, dtype=torch.float64, device=dev)
        rng = np.random.default_rng(11)
        for it in range(7):                                   # both images, step numbers 1..7, x new every step
            x = rng.standard_normal(blk.m)
            xd = torch.from_numpy(x).to(dev)
            nxt = h.lib.rsg_dist_x(d)
            target = img[0] if nxt == img[0].data_ptr() else img[1]
            assert nxt in (img[0].data_ptr(), img[1].data_ptr())
            target[:blk.m].copy_(xd)
            h.dist_step(d, y)
            h.synchronize()
            assert np.array_equal(y.cpu().numpy(), _loopback_reference(blk, hbase, send, x)), (proto, it)
        for it in range(40):                                  # back to back, no host sync in between
            h.dist_step(d, y)
        h.synchronize()
        assert np.array_equal(y.cpu().numpy(), _loopback_reference(blk, hbase, send, x)), proto
        assert h.info("SPMV_ALGO_USED") == 8
        h.dist_destroy(d)
    finally:
        h.close()


def test_single_device_packets_that_never_arrive_time_out():
    """the product's halo tiles poll packets nobody sends (the push goes to another buffer): RSG_DEPENDENCY_TIMEOUT, no hang"""
    from respasol_b200.gpu import DeviceCSR, Handle, RsgError
    dev = torch.device("cuda", 0)
    blk, hbase, send, plane = _loopback_problem(nx=16, nz=8)
    h = Handle(0)
    try:
        A = DeviceCSR(blk, "d", dev)
        stride = (blk.n + 31) // 32 * 32
        ximg = torch.zeros(2 * stride, dtype=torch.float64, device=dev)
        flags = torch.zeros(128, dtype=torch.int32, device=dev)
        send_d = torch.from_numpy(send).to(dev)
        img = [ximg[p * stride:] for p in range(2)]
        d = h.dist_create(A, [t.data_ptr() for t in img], hbase, send_d, [[img[p].data_ptr() + 8 * hbase] for p in range(2)],
                          [[flags[32 * p:].data_ptr()] for p in range(2)], [0], [len(send)], [flags[32 * p:].data_ptr() for p in range(2)], 1)
        pk = torch.zeros(2 * 2 * plane * 2, dtype=torch.float64, device=dev)
        elsewhere = torch.zeros(2 * 2 * plane * 2, dtype=torch.float64, device=dev)
        mine = [pk.data_ptr() + 16 * 2 * plane * p for p in range(2)]
        lost = [elsewhere.data_ptr() + 16 * 2 * plane * p for p in range(2)]
        assert h.dist_enable_packets(d, mine, [[lost[0]], [lost[1]]], side_stream=True)
        h.set_option("WAIT_SPINS", 2000)
        y = torch.zeros(blk.m, dtype=torch.float64, device=dev)
        h.dist_step(d, y)
        with pytest.raises(RsgError) as ei:
            h.synchronize()
        assert ei.value.status == 7
        h.set_option("WAIT_SPINS", 0)
        h.dist_destroy(d)
    finally:
        h.close()
