#!/usr/bin/env python
"""Small cases of every flag- / barrier-synchronised kernel, for compute-sanitizer (one tool per run):
    compute-sanitizer --tool memcheck|racecheck|synccheck python tools/sanitize_cases.py
SpMV window / tile / vector / merge (ragged ends, long-row segments, halo push + in-kernel wait emulated in ONE stream),
device-built level sets, ILU0 (sync-free flags), L/U SpTRSV (sentinel polling and level-scheduled), BiCGStab, transpose.
Every result is checked against the oracle, so a sanitizer-clean run is also a correct one."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402
import torch  # noqa: E402

import orc  # noqa: E402
from respasol_b200 import host  # noqa: E402
from respasol_b200.dist import HaloPlan  # noqa: E402
from respasol_b200.gpu import ALGO, DeviceCSR, Handle  # noqa: E402

h = Handle(0)
n_checked = 0


def spmv_case(a, algo, precision="d", **opts):
    global n_checked
    h.set_option("SPMV_ALGO", ALGO[algo])
    for k, v in opts.items():
        h.set_option(k, v)
    x = orc.dlarnv(a.n) - 0.5
    A = DeviceCSR(a, precision)
    xt = torch.float32 if precision == "s" else torch.float64
    y = h.spmv(A, torch.from_numpy(x).cuda().to(xt)).cpu().numpy().astype(np.float64)
    ref = orc.spmv(a, x, precision).astype(np.float64)
    assert np.abs(y - ref).max() <= (1e-12 if precision == "d" else 5e-5) * max(np.abs(ref).max(), 1e-300), (algo, precision)
    for k in opts:
        h.set_option(k, 0)
    h.set_option("SPMV_ALGO", 0)
    n_checked += 1


mats = [host.gen_stencil27(9, 10, 11, seed=5), host.gen_laplace2d(23), host.gen_femlike(8, 0.7, 9),
        host.gen_powerlaw(3000, 2.2, 48, 2, 2600, 4)]
for a in mats:
    for p in ("d", "s", "m"):
        for algo in ("window", "tile", "vector", "merge", "rowblock"):
            spmv_case(a, algo, p)
spmv_case(mats[0], "window", "d", SPMV_WIN_BYTES=8192, SPMV_VARIANT=1)
spmv_case(mats[0], "window", "d", SPMV_REDUCE=8)
spmv_case(mats[3], "tile", "d", SPMV_LONG_ROW=64)

# halo push + in-kernel wait, two ranks emulated in one stream (window and tile kernels, fused and separate push)
full = host.gen_stencil27(12, 12, 20, seed=2)
starts = np.array([0, full.m // 2, full.m])
plans = [HaloPlan(full.rows(int(starts[r]), int(starts[r + 1])), starts, r) for r in range(2)]
needs = [p.need for p in plans]
for p in plans:
    p.finish(needs)
p0, p1 = plans
x = orc.dlarnv(full.n) - 0.5
ref = orc.spmv(full, x, "d")[:p0.n_local]
for algo in ("window", "tile"):
    h.set_option("SPMV_ALGO", ALGO[algo])
    h.set_option("SPMV_HALO_START", p0.halo_base)
    A = DeviceCSR(p0.remapped, "d")
    x_ext = torch.zeros(p0.n_ext, dtype=torch.float64, device="cuda")
    x_ext[:p0.n_local] = torch.from_numpy(x[:p0.n_local]).cuda()
    x1 = torch.from_numpy(x[p0.n_local:]).cuda()
    send1 = torch.from_numpy(p1.my_send.astype(np.int32)).cuda()
    send0 = torch.from_numpy(p0.my_send.astype(np.int32)).cuda()
    flags = torch.zeros(32, dtype=torch.int32, device="cuda")
    peer_halo = torch.zeros(p1.n_ext, dtype=torch.float64, device="cuda")
    peer_flags = torch.zeros(32, dtype=torch.int32, device="cuda")
    y = torch.empty(p0.n_local, dtype=torch.float64, device="cuda")
    for epoch in (1, 2):
        flags[0] = epoch
        h.halo_push(send1, x1, [x_ext[p0.halo_base + p0.width:].data_ptr()], [flags[1:].data_ptr()], epoch)
        h.spmv_halo_fuse(send0, [0, peer_halo[p1.halo_base:].data_ptr()], [0, peer_flags.data_ptr()], epoch)
        h.spmv_halo_wait(flags, 2, epoch)
        h.spmv(A, x_ext, y)
        h.synchronize()
        assert np.abs(y.cpu().numpy() - ref).max() <= 1e-12 * np.abs(ref).max(), algo
        assert int(peer_flags[0].item()) == epoch
        n_checked += 1
    h.set_option("SPMV_HALO_START", 0)
    h.set_option("SPMV_ALGO", 0)

# level sets, ILU0, SpTRSV (sync-free and level-scheduled), BiCGStab, transpose
for a in (host.gen_laplace2d(17), host.gen_femlike(9, 0.6, 3)):
    for dt, p in ((np.float64, "d"), (np.float32, "s")):
        A = DeviceCSR(a, p)
        h.ilu0_analysis(A)
        lp, rows = h.levels(0, a.m)
        _, lp_ref, rows_ref = orc.levels(a, 0)
        assert np.array_equal(lp, lp_ref) and np.array_equal(rows, rows_ref)
        st, piv = h.ilu0_factor(A)
        _, _, lu_ref = orc.ilu0(a, dt)
        assert st == 0 and np.array_equal(A.values.cpu().numpy(), lu_ref)
        b = torch.ones(a.m, dtype=A.values.dtype, device="cuda")
        for algo in (0, 2):
            h.set_option("TRSV_ALGO", algo)
            z, yv = h.sptrsv_lu(A, A.values, b)
            _, _, y_ref = orc.sptrsv_lu(a, lu_ref, np.ones(a.m, dtype=dt))
            assert np.array_equal(yv.cpu().numpy(), y_ref), ("trsv", p, algo)
            n_checked += 1
        h.set_option("TRSV_ALGO", 0)
f = host.gen_femlike(9, 0.6, 3)
A, F = DeviceCSR(f, "d"), DeviceCSR(f, "d")
h.ilu0_factor(F)
xt = torch.from_numpy(orc.dlarnv(f.m) - 0.5).cuda()
rhs = h.spmv(A, xt).clone()
for solver in ("bicgstab", "gmres", "pcg"):
    xk, it, rel = getattr(h, solver)(A, rhs, lu=F.values, max_iters=60, rtol=1e-9)
    n_checked += 1
assert float((h.bicgstab(A, rhs, lu=F.values, max_iters=60, rtol=1e-10)[0] - xt).abs().max()) < 1e-6
yt = h.spmv_t(A, xt)
assert abs(float((yt * torch.ones_like(yt)).sum()) - float((xt * h.spmv(A, torch.ones_like(xt))).sum())) < 1e-9
h.synchronize()
print("sanitize_cases ok: %d checked cases, %d kernel launches" % (n_checked, h.info("KERNEL_LAUNCHES")))
h.close()
