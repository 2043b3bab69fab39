#!/usr/bin/env python
"""A/B one handle option on the SpMV configs, cold L2 (rotated replicas), interleaved repetitions.
    python tools/ab_option.py SPMV_X_HINT 0 2 [--configs C2,C5] [--prec d,s] [--reps 300]"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import numpy as np  # noqa: E402
import torch  # noqa: E402
import orc  # noqa: E402
from respasol_b200.gpu import DeviceCSR, Handle  # noqa: E402
from spmv_sweep import matrix, spmv_bytes  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("option")
ap.add_argument("a", type=int)
ap.add_argument("b", type=int)
ap.add_argument("--configs", default="C2")
ap.add_argument("--prec", default="d,s")
ap.add_argument("--reps", type=int, default=300)
ap.add_argument("--set", default="", help="fixed options for both arms: NAME=VAL,NAME=VAL")
args = ap.parse_args()
h = Handle(0)
for kv in filter(None, args.set.split(",")):
    h.set_option(kv.split("=")[0], int(kv.split("=")[1]))
for cname in args.configs.split(","):
    a = matrix(cname)
    x = orc.dlarnv(a.n) - 0.5
    for p in args.prec.split(","):
        b = spmv_bytes(a.stored, a.m, a.n, p)
        R = max(2, int(np.ceil(300e6 / b)) + 1)
        xt = torch.float32 if p == "s" else torch.float64
        mats = [DeviceCSR(a, p) for _ in range(R)]
        xs = torch.from_numpy(x).cuda().to(xt)
        y = torch.empty(a.m, dtype=xt, device="cuda")
        res = {args.a: [], args.b: []}
        for rnd in range(5):
            for val in (args.a, args.b):
                h.set_option(args.option, val)
                for r in range(R):
                    h.spmv(mats[r], xs, y)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for i in range(args.reps):
                    h.spmv(mats[i % R], xs, y)
                e1.record()
                torch.cuda.synchronize()
                res[val].append(e0.elapsed_time(e1) / args.reps * 1e3)
        for val in (args.a, args.b):
            us = float(np.median(res[val]))
            print("%s %s %s=%d: %.2f us  %.0f GB/s  (runs %s)" % (cname, p, args.option, val, us, b / us * 1e-3,
                                                                 " ".join("%.2f" % t for t in res[val])), flush=True)
        del mats
        torch.cuda.empty_cache()
