#!/usr/bin/env python
"""Run a handful of launches of one SpMV kernel configuration (for ncu):
    python tools/prof_one.py C2 d stream 2 [xhint] [n_launches]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch  # noqa: E402

import orc  # noqa: E402
from respasol_b200.gpu import ALGO, DeviceCSR, Handle  # noqa: E402
sys.path.insert(0, os.path.join(ROOT, "tools"))
from spmv_sweep import matrix  # noqa: E402

cfg, prec, algo, variant = sys.argv[1], sys.argv[2], sys.argv[3], int(sys.argv[4])
xh = int(sys.argv[5]) if len(sys.argv) > 5 else 0
n = int(sys.argv[6]) if len(sys.argv) > 6 else 6
a = matrix(cfg)
h = Handle(0)
h.set_option("SPMV_ALGO", ALGO[algo])
h.set_option("SPMV_VARIANT", variant)
h.set_option("SPMV_X_HINT", xh)
for kv in os.environ.get("RSG_OPTS", "").split(","):
    if "=" in kv:
        k, v = kv.split("=")
        h.set_option(k, int(v))
xt = torch.float32 if prec == "s" else torch.float64
mats = [DeviceCSR(a, prec) for _ in range(2)]
x = torch.from_numpy(orc.dlarnv(a.n) - 0.5).cuda().to(xt)
y = torch.empty(a.m, dtype=xt, device="cuda")
for i in range(n):
    h.spmv(mats[i % 2], x, y)
torch.cuda.synchronize()
print("done", float(y.abs().sum()))
