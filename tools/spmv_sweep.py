#!/usr/bin/env python
"""Sweep SpMV kernels / shapes / hints on the BASELINE configs (GPU box).  Cold-L2 timing: R replicas
of the matrix are rotated so that consecutive products never find their stream in the 126 MB L2.
Every timed combination is first checked against the plain sub-warp-per-row kernel.
    python tools/spmv_sweep.py [--configs C1,C2,C4,C5] [--prec d,s,m] [--reps 200] [--out gpurun_out/sweep.jsonl]
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
from respasol_b200.gpu import ALGO, DeviceCSR, Handle  # noqa: E402


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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--configs", default="C1,C2,C4")
    ap.add_argument("--prec", default="d,s,m")
    ap.add_argument("--reps", type=int, default=200)
    ap.add_argument("--combos", default="all")
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "sweep.jsonl"))
    args = ap.parse_args()
    h = Handle(0)
    dev = torch.device("cuda", 0)
    out = open(args.out, "a")
    combos = []
    for v in range(8):
        for lanes in (0,):  # pipe: the "xhint" slot carries the forced lanes per row
            combos.append(("pipe", v, lanes, 1))
    for v in range(10):
        for ts in (0,):  # tile: the "xhint" slot carries the tile size (0 = heuristic)
            combos.append(("tile", v, ts, 1))
    for v in range(6):
        for xh in (0, 1):
            combos.append(("stream", v, xh, 1))
    for v in range(5):
        combos.append(("rowblock", v, 0, 1))
    combos.append(("rowblock", 0, 0, 2))
    combos.append(("merge", 0, 0, 1))
    combos.append(("vector", 0, 0, 1))
    if args.combos != "all":
        want = set(args.combos.split(","))
        combos = [c for c in combos if "%s%d%d" % (c[0], c[1], c[2]) in want or c[0] in want]
    for cname in args.configs.split(","):
        a = matrix(cname)
        x = orc.dlarnv(a.n) - 0.5
        for p in args.prec.split(","):
            b = spmv_bytes(a.stored, a.m, a.n, p)
            R = max(2, int(np.ceil(300e6 / b)) + 1)  # rotate > 300 MB of distinct streams
            xt = torch.float32 if p == "s" else torch.float64
            mats = [DeviceCSR(a, p, dev) for _ in range(R)]
            xs = [torch.from_numpy(x).to(dev).to(xt) for _ in range(R)]
            ys = [torch.empty(a.m, dtype=xt, device=dev) for _ in range(R)]
            h.set_option("SPMV_ALGO", ALGO["vector"])
            yref = h.spmv(mats[0], xs[0]).clone()
            scale = float(yref.abs().max())
            for algo, var, xh, hint in combos:
                h.set_option("SPMV_ALGO", ALGO[algo])
                h.set_option("SPMV_VARIANT", var)
                h.set_option("SPMV_X_HINT", xh if algo not in ("pipe", "tile") else 0)
                h.set_option("SPMV_LONG_ROW", xh if algo == "tile" else 0)
                h.set_option("SPMV_REDUCE", xh if algo == "pipe" else 0)
                h.set_option("SPMV_CACHE_HINT", hint if algo != "pipe" else 1)
                h.set_option("SPMV_ROWS_PER_THREAD", hint if algo == "pipe" else 0)
                try:
                    for r in range(R):
                        h.spmv(mats[r], xs[r], ys[r])
                    torch.cuda.synchronize()
                    err = float((ys[0] - yref).abs().max()) / max(scale, 1e-300)
                    ok = err <= (1e-12 if p != "s" else 1e-4)
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record()
                    for i in range(args.reps):
                        r = i % R
                        h.spmv(mats[r], xs[r], ys[r])
                    e1.record()
                    torch.cuda.synchronize()
                    us = e0.elapsed_time(e1) / args.reps * 1e3
                    rec = {"config": cname, "prec": p, "algo": algo, "variant": var, "xhint": xh, "hint": hint,
                           "us": round(us, 2), "GBps": round(b / us * 1e-3, 1), "ok": ok, "relerr": err,
                           "used": h.info("SPMV_ALGO_USED"), "blocks": h.info("SPMV_ROW_BLOCKS"), "replicas": R}
                except Exception as e:  # noqa: BLE001
                    rec = {"config": cname, "prec": p, "algo": algo, "variant": var, "xhint": xh, "error": repr(e)}
                print(json.dumps(rec), flush=True)
                out.write(json.dumps(rec) + "\n")
                out.flush()
            del mats, xs, ys
            torch.cuda.empty_cache()
    h.set_option("SPMV_ALGO", 0)
    h.set_option("SPMV_ROWS_PER_THREAD", 0)
    h.set_option("SPMV_REDUCE", 0)
    h.set_option("SPMV_LONG_ROW", 0)
    h.close()


if __name__ == "__main__":
    main()
