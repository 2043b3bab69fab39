#!/usr/bin/env python
"""Sweep SpMV kernels / shapes / options on the BASELINE configs (GPU box).  Cold-L2 timing: R replicas
of the matrix are rotated so that consecutive products never find their stream in L2 (> 300 MB of distinct streams).
Every timed combination is first checked against the plain sub-warp-per-row kernel.
    python tools/spmv_sweep.py [--configs C1,C2,C3,C4,C5] [--prec d,s,m] [--reps 200] [--combos SPEC;SPEC...]
A SPEC is  algo[:OPTION=value[,OPTION=value...]]  e.g.  window:SPMV_WIN_BYTES=36864,SPMV_VARIANT=1
Default combos: the window kernel over tile budgets / CTA widths, the tile kernel with and without the residency
hint (variants 0/1 vs 6/7), the batched sub-warp kernel, AUTO.
"""
import argparse
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
from respasol_b200.gpu import ALGO, OPT, DeviceCSR, Handle  # noqa: E402


def matrix(name):
    if name == "C1":
        return host.gen_laplace2d(1000)
    if name == "C2":
        return host.gen_stencil27(100, seed=2)
    if name == "C3":
        return host.gen_femlike(47, 0.58, 3)
    if name == "C4":
        return host.gen_powerlaw(320000, 2.2, 64, 4, 100000, 4)
    if name == "C5":
        return host.gen_stencil27(134, seed=2)
    raise KeyError(name)


def spmv_bytes(stored, m, n, p):
    v, x = {"d": (8, 8), "s": (4, 4), "m": (4, 8)}[p]
    return stored * (v + 4) + (m + 1) * 4 + m * x + n * x


DEFAULT = ["auto", "vector"]
for wb in (28672, 36864, 45056, 56320, 75776):
    DEFAULT.append("window:SPMV_WIN_BYTES=%d" % wb)
for pct in (56, 72):
    DEFAULT.append("window:SPMV_WIN_STREAM_PCT=%d" % pct)
DEFAULT += ["window:SPMV_PDL=1", "window:SPMV_PDL=2", "window:SPMV_PDL=2,SPMV_WIN_BYTES=28672", "window:SPMV_PDL=2,SPMV_WIN_BYTES=45056",
            "window:SPMV_PDL=2,SPMV_VARIANT=1,SPMV_WIN_BYTES=22528", "window:SPMV_PDL=2,SPMV_VARIANT=2,SPMV_WIN_BYTES=75776"]
DEFAULT += ["window:SPMV_VARIANT=1,SPMV_WIN_BYTES=22528", "window:SPMV_VARIANT=1,SPMV_WIN_BYTES=28672",
            "window:SPMV_VARIANT=2,SPMV_WIN_BYTES=75776", "window:SPMV_VARIANT=2,SPMV_WIN_BYTES=110592",
            "window:SPMV_REDUCE=2", "window:SPMV_REDUCE=4",
            "tile:SPMV_VARIANT=0", "tile:SPMV_VARIANT=1", "tile:SPMV_VARIANT=5", "tile:SPMV_VARIANT=6", "tile:SPMV_VARIANT=7"]


def parse(spec):
    algo, _, rest = spec.partition(":")
    opts = {}
    for kv in filter(None, rest.split(",")):
        k, v = kv.split("=")
        opts[k] = int(v)
    return algo, opts


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--configs", default="C1,C2,C4")
    ap.add_argument("--prec", default="d,s,m")
    ap.add_argument("--reps", type=int, default=200)
    ap.add_argument("--combos", default="")
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "sweep.jsonl"))
    args = ap.parse_args()
    h = Handle(0)
    dev = torch.device("cuda", 0)
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    out = open(args.out, "a")
    combos = [parse(s) for s in (args.combos.split(";") if args.combos else DEFAULT)]
    for cname in args.configs.split(","):
        a = matrix(cname)
        x = orc.dlarnv(a.n) - 0.5
        for p in args.prec.split(","):
            b = spmv_bytes(a.stored, a.m, a.n, p)
            R = max(2, int(np.ceil(300e6 / b)) + 1)
            xt = torch.float32 if p == "s" else torch.float64
            mats = [DeviceCSR(a, p, dev) for _ in range(R)]
            xs = [torch.from_numpy(x).to(dev).to(xt) for _ in range(R)]
            ys = [torch.empty(a.m, dtype=xt, device=dev) for _ in range(R)]
            h.set_option("SPMV_ALGO", ALGO["vector"])
            yref = h.spmv(mats[0], xs[0]).clone()
            scale = float(yref.abs().max())
            for algo, opts in combos:
                h.set_option("SPMV_ALGO", ALGO[algo])
                for k, v in opts.items():
                    h.set_option(k, v)
                try:
                    for r in range(R):
                        h.spmv(mats[r], xs[r], ys[r])
                    torch.cuda.synchronize()
                    err = float((ys[0] - yref).abs().max()) / max(scale, 1e-300)
                    ok = err <= (1e-12 if p != "s" else 1e-4)
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    best = None
                    for _ in range(3):  # three rounds, the best is kept (the spread is printed)
                        e0.record()
                        for i in range(args.reps):
                            r = i % R
                            h.spmv(mats[r], xs[r], ys[r])
                        e1.record()
                        torch.cuda.synchronize()
                        us = e0.elapsed_time(e1) / args.reps * 1e3
                        best = us if best is None else min(best, us)
                    rec = {"config": cname, "prec": p, "algo": algo, "opts": opts, "us": round(best, 2), "us_last": round(us, 2),
                           "GBps": round(b / best * 1e-3, 1), "ok": ok, "relerr": err, "used": h.info("SPMV_ALGO_USED"),
                           "blocks": h.info("SPMV_ROW_BLOCKS"), "workspace": h.info("WORKSPACE_BYTES"), "replicas": R}
                except Exception as e:  # noqa: BLE001
                    rec = {"config": cname, "prec": p, "algo": algo, "opts": opts, "error": repr(e)}
                for k in opts:
                    h.set_option(k, 0)
                print(json.dumps(rec), flush=True)
                out.write(json.dumps(rec) + "\n")
                out.flush()
            del mats, xs, ys
            torch.cuda.empty_cache()
    h.set_option("SPMV_ALGO", 0)
    h.close()


if __name__ == "__main__":
    main()
