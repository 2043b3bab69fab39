#!/usr/bin/env python
"""L+U SpTRSV (and ILU0 factor) milliseconds per algorithm on config 3, the thin-level 2-D 5-point 1000^2 and a 27-point 64^3.
    python tools/trsv_ab.py"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402
import torch  # noqa: E402

import orc  # noqa: E402
from respasol_b200 import host  # noqa: E402
from respasol_b200.gpu import DeviceCSR, Handle  # noqa: E402

h = Handle(0)
cases = (("c3_femlike47", host.gen_femlike(47, 0.58, 3)), ("lap2d_1000", host.gen_laplace2d(1000)), ("stencil27_64", host.gen_stencil27(64, seed=2)),
         ("lap2d_200", host.gen_laplace2d(200)))
for name, a in cases:
    for p, dt in (("d", torch.float64), ("s", torch.float32)):
        A = DeviceCSR(a, p)
        h.ilu0_factor(A, report_pivot=False)
        b = torch.ones(a.m, dtype=dt, device="cuda")
        z, y = torch.empty_like(b), torch.empty_like(b)
        ref = None
        row = {"case": name, "prec": p, "n": a.m, "stored": a.stored}
        for algo in (1, 3, 2, 0):
            h.set_option("TRSV_ALGO", algo)
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
            ts = []
            for it in range(8):
                torch.cuda.synchronize()
                ev[0].record(); h.sptrsv_lu(A, A.values, b, z, y); ev[1].record(); torch.cuda.synchronize()
                ts.append(ev[0].elapsed_time(ev[1]))
            h.synchronize()
            if ref is None:
                ref = y.clone()
            row["algo%d_ms" % algo] = round(float(np.median(ts[2:])), 4)
            row["algo%d_same_bits" % algo] = bool(torch.equal(y, ref))
        h.set_option("TRSV_ALGO", 0)
        row["levels"] = (h.info("LEVELS_L"), h.info("LEVELS_U"))
        print(json.dumps(row), flush=True)
h.close()
