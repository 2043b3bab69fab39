#!/usr/bin/env python
"""Per-CTA timeline (rsg_debug_trace) of back-to-back single-GPU products: python tools/tile_trace.py [--config C4] [--prec d] [--opts K=V,...]
For the LAST product of a cold-L2 run (replicas rotated): when the product before it ended, when this one's launch dependency
resolved, and entry / staged / done percentiles of its CTAs by kind (tiles; long-row segment CTAs)."""
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
from respasol_b200.gpu import ALGO, DeviceCSR, Handle  # noqa: E402
from spmv_sweep import matrix, spmv_bytes  # noqa: E402


def pct(a, qs=(0, 10, 50, 90, 100)):
    return [round(float(np.percentile(a, q)), 2) for q in qs] if len(a) else []


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="C4")
    ap.add_argument("--prec", default="d")
    ap.add_argument("--algo", default="auto")
    ap.add_argument("--opts", default="SPMV_PDL=2")
    ap.add_argument("--steps", type=int, default=301)
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    h = Handle(0)
    h.set_option("SPMV_ALGO", ALGO[args.algo])
    for kv in filter(None, args.opts.split(",")):
        k, v = kv.split("=")
        h.set_option(k, int(v))
    a = matrix(args.config)
    x = orc.dlarnv(a.n) - 0.5
    b = spmv_bytes(a.stored, a.m, a.n, args.prec)
    R = max(2, int(np.ceil(300e6 / b)) + 1)
    xt = torch.float32 if args.prec == "s" else torch.float64
    mats = [DeviceCSR(a, args.prec, dev) for _ in range(R)]
    xs = [torch.from_numpy(x).to(dev).to(xt) for _ in range(R)]
    ys = [torch.empty(a.m, dtype=xt, device=dev) for _ in range(R)]
    for r in range(R):
        h.spmv(mats[r], xs[r], ys[r])
    torch.cuda.synchronize()
    cap = 1 << 17
    buf = torch.zeros(2 * 4 * cap, dtype=torch.int64, device=dev)
    assert h.lib.rsg_debug_trace(h.h, buf.data_ptr(), cap) == 0
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):
        r = i % R
        h.spmv(mats[r], xs[r], ys[r])
    e1.record()
    torch.cuda.synchronize()
    assert h.lib.rsg_debug_trace(h.h, None, 0) == 0
    us = e0.elapsed_time(e1) / args.steps * 1e3
    t = buf.cpu().numpy().astype(np.uint64).reshape(2, cap, 4)
    last = (args.steps - 1) & 1
    cur, prev = t[last], t[last ^ 1]
    seg = (cur[:, 3] >> np.uint64(63)).astype(bool)
    mask = np.uint64((1 << 62) - 1)
    cur = (cur & mask).astype(np.int64)
    prev = (prev & mask).astype(np.int64)
    used = cur[:, 0] > 0
    n = int(np.flatnonzero(used)[-1]) + 1 if used.any() else 0
    cur, seg, used = cur[:n], seg[:n], used[:n]
    T0 = cur[used, 0].min()
    rel = (cur - T0) / 1e3
    out = {"config": args.config, "prec": args.prec, "opts": args.opts, "us_per_product": round(us, 2), "algo": h.info("SPMV_ALGO_USED"),
           "ctas": n, "segment_ctas": int(seg.sum()), "prev_product_end": round(float((prev[:, 3].max() - T0) / 1e3), 2),
           "this_product_end": round(float(rel[used, 3].max()), 2)}
    for name, sel in (("tiles", used & ~seg & (cur[:, 3] > 0)), ("segments", used & seg)):
        v = rel[sel]
        if len(v):
            out[name] = {"n": int(sel.sum()), "entry": pct(v[:, 0]), "dependency": pct(v[:, 1]), "staged": pct(v[:, 2]), "done": pct(v[:, 3]),
                         "staged_minus_dep": pct(v[:, 2] - np.maximum(v[:, 1], v[:, 0])), "done_minus_staged": pct(v[:, 3] - v[:, 2]),
                         "life": pct(v[:, 3] - v[:, 0])}
    print(json.dumps(out), flush=True)
    h.close()


if __name__ == "__main__":
    main()
