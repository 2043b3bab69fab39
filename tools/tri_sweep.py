#!/usr/bin/env python
"""Sweep of the sync-free ILU0 / SpTRSV launch shapes on config 3 (FEM-like, ~104k rows, 321 levels).
    python tools/tri_sweep.py [--grid 47] [--reps 10]
Prints one line per combination: batch, gate, CTAs/SM cap, block -> L+U solve ms (median), and the
factorization ms per CTA cap.  Every variant is checked bit-exact against the first one."""
import argparse
import itertools
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402
from respasol_b200 import host  # noqa: E402
from respasol_b200.gpu import DeviceCSR, Handle  # noqa: E402


def med(fn, reps):
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    ts = []
    for _ in range(reps + 1):
        torch.cuda.synchronize()
        ev[0].record()
        fn()
        ev[1].record()
        torch.cuda.synchronize()
        ts.append(ev[0].elapsed_time(ev[1]))
    return float(np.median(ts[1:])), float(np.min(ts[1:]))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--grid", type=int, default=47)
    ap.add_argument("--reps", type=int, default=10)
    ap.add_argument("--precision", default="d")
    args = ap.parse_args()
    h = Handle(0)
    a = host.gen_femlike(args.grid, 0.58, 3)
    A = DeviceCSR(a, args.precision)
    dt = torch.float64 if args.precision == "d" else torch.float32
    pristine = A.values.clone()
    h.ilu0_analysis(A)
    h.sptrsv_analysis(A)
    print("n=%d stored=%d levels L/U = %d/%d" % (a.m, a.stored, h.info("LEVELS_L"), h.info("LEVELS_U")))
    # factorization vs CTA cap
    lu_first = None
    for cap in (0, 1, 2, 3, 4):
        h.set_option("TRI_CTAS_PER_SM", cap)

        def fac():
            A.values.copy_(pristine)
            h.ilu0_factor(A, report_pivot=False)
        fac()
        tcopy, _ = med(lambda: A.values.copy_(pristine), args.reps)
        t, tmin = med(fac, args.reps)
        lu = A.values.clone()
        if lu_first is None:
            lu_first = lu
        same = bool(torch.equal(lu, lu_first))
        print("ilu0 cap=%d  factor %.3f ms (min %.3f, copy %.3f subtracted)  same=%s" % (cap, t - tcopy, tmin - tcopy, tcopy, same))
    h.set_option("TRI_CTAS_PER_SM", 0)
    A.values.copy_(lu_first)
    b = torch.ones(a.m, dtype=dt, device="cuda")
    z, y = torch.empty_like(b), torch.empty_like(b)
    y_first = None
    rows = []
    for gate, batch, cap, block in itertools.product((1, 2), (4, 8, 16), (0, 1, 2), (32, 64, 128)):
        h.set_option("TRSV_GATE", gate)
        h.set_option("TRSV_BATCH", batch)
        h.set_option("TRI_CTAS_PER_SM", cap)
        h.set_option("TRSV_BLOCK", block)
        t, tmin = med(lambda: h.sptrsv_lu(A, A.values, b, z, y), args.reps)
        if y_first is None:
            y_first = y.clone()
        same = bool(torch.equal(y, y_first))
        rows.append((t, gate, batch, cap, block, tmin, same))
        print("trsv gate=%d batch=%d cap=%d block=%3d  L+U %.3f ms (min %.3f) same=%s" % (gate, batch, cap, block, t, tmin, same))
    rows.sort()
    print("best:", rows[:5])
    h.set_option("TRSV_ALGO", 2)
    for k in ("TRSV_GATE", "TRSV_BATCH", "TRI_CTAS_PER_SM", "TRSV_BLOCK"):
        h.set_option(k, 0)
    t, tmin = med(lambda: h.sptrsv_lu(A, A.values, b, z, y), args.reps)
    print("level-scheduled launches: L+U %.3f ms (min %.3f) same=%s" % (t, tmin, bool(torch.equal(y, y_first))))
    h.close()


if __name__ == "__main__":
    main()
