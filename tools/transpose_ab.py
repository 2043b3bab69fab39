#!/usr/bin/env python
"""A/B of the transposed product's three value modes and the fused kernel's lanes per column, cold L2.
    python tools/transpose_ab.py > gpurun_out/transpose_ab.json"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402
from respasol_b200 import host  # noqa: E402
from respasol_b200.gpu import DeviceCSR, Handle  # noqa: E402


def timed(fn, reps, flush):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for _ in range(5):
        fn()
    ts = []
    for i in range(reps):
        flush.fill_(i & 0xff)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    return float(np.median(ts))


def main():
    h = Handle(0)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    out = {}
    for name, a in (("c1_lap2d_1000", host.gen_laplace2d(1000)), ("c2_stencil27_100", host.gen_stencil27(100, seed=2)),
                    ("c5_stencil27_134", host.gen_stencil27(134, seed=2)), ("c3_femlike_47", host.gen_femlike(47, 0.58, 3))):
        for prec in ("d", "s"):
            A = DeviceCSR(a, prec)
            x = torch.rand(a.m, dtype=A.values.dtype, device="cuda") - 0.5
            y = torch.empty(a.n, dtype=A.values.dtype, device="cuda")
            row = {"A_us": timed(lambda: h.spmv(A, x[:a.n] if a.n <= a.m else x, torch.empty(a.m, dtype=x.dtype, device="cuda")), 20, flush)
                   if a.m == a.n else None}
            for mode, mn in ((2, "gather+tile"), (1, "static"), (0, "fused_auto")):
                h.set_option("TRANSPOSE_STATIC", mode)
                row[mn + "_us"] = timed(lambda: h.spmv_t(A, x, y), 20, flush)
            for lanes in (1, 2, 4, 8, 16):
                h.set_option("SPMV_REDUCE", lanes)
                row["fused_%d_lanes_us" % lanes] = timed(lambda: h.spmv_t(A, x, y), 20, flush)
            h.set_option("SPMV_REDUCE", 0)
            out[name + "_" + prec] = row
            print(name, prec, json.dumps(row), file=sys.stderr, flush=True)
            h.lib.rsg_forget_matrix(h.h, A.rowptr.data_ptr())
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
