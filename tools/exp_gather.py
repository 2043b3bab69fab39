#!/usr/bin/env python
"""Experiment: how much of the SpMV time is the x gather?  Times kernels on config C2 with the real
column indices, with colidx := own row (gathers always hit L1) and with random columns."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import numpy as np  # noqa: E402
import torch  # noqa: E402

from respasol_b200.gpu import ALGO, DeviceCSR, Handle  # noqa: E402
from spmv_sweep import matrix  # noqa: E402

cfg = sys.argv[1] if len(sys.argv) > 1 else "C2"
a = matrix(cfg)
h = Handle(0)
R = 3
rows = np.repeat(np.arange(a.m, dtype=np.int32), np.diff(a.rowptr))
rng = np.random.default_rng(1)
cols = {"real": a.colidx, "own_row": rows, "random": np.sort(rng.integers(0, a.n, size=(a.m, 1), dtype=np.int32).repeat(1, 1), axis=1).ravel()[rows] if False else rng.integers(0, a.n, size=a.stored, dtype=np.int32)}
# ceiling: plain streaming reads of the same number of bytes (torch.sum = CUB reduce) and a copy
for nbytes in (340e6, 680e6):
    n = int(nbytes // 8)
    bufs = [torch.ones(n, dtype=torch.float64, device="cuda") for _ in range(3)]
    for name, fn, moved in (("sum(read)", lambda b: b.sum(), 1), ("copy(r+w)", lambda b: bufs[0].copy_(b) if b is not bufs[0] else bufs[1].copy_(b), 2)):
        for b in bufs:
            fn(b)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(60):
            fn(bufs[i % 3])
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) / 60 * 1e3
        print("ceiling %s %.0f MB: %.2f us  %.0f GB/s" % (name, nbytes / 1e6, us, moved * n * 8 / us * 1e-3), flush=True)
    del bufs
    torch.cuda.empty_cache()
for prec in ("d", "s"):
    xt = torch.float32 if prec == "s" else torch.float64
    x = torch.ones(a.n, dtype=xt, device="cuda")
    for cname, c in cols.items():
        mats = []
        for _ in range(R):
            A = DeviceCSR(a, prec)
            A.colidx = torch.from_numpy(np.ascontiguousarray(c)).cuda()
            mats.append(A)
        y = torch.empty(a.m, dtype=xt, device="cuda")
        for algo, var in (("tile", 0), ("tile", 1), ("tile", 2), ("stream", 2), ("stream", 0), ("pipe", 3)):
            h.set_option("SPMV_ALGO", ALGO[algo])
            h.set_option("SPMV_VARIANT", var)
            for A in mats:
                h.spmv(A, x, y)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for i in range(150):
                h.spmv(mats[i % R], x, y)
            e1.record()
            torch.cuda.synchronize()
            print("%s %s cols=%-8s %-8s v%d  %.2f us" % (cfg, prec, cname, algo, var, e0.elapsed_time(e1) / 150 * 1e3), flush=True)
        del mats
        torch.cuda.empty_cache()
