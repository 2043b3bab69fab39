#!/usr/bin/env python
"""Where the host-vector product (rsg_matrix_spmv_host, bench.py's e2e leg) spends its time on config 2:
PCIe copies alone / together, and the pipelined call for several chunk counts.
    python tools/e2e_breakdown.py [--reps 50]"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402
from respasol_b200 import host  # noqa: E402
from respasol_b200.gpu import Handle  # noqa: E402


def wall(fn, reps):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps * 1e6


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reps", type=int, default=50)
    ap.add_argument("--grid", type=int, default=100)
    args = ap.parse_args()
    a = host.gen_stencil27(args.grid, seed=2)
    n = a.n
    out = {"bytes_each_way": n * 8}
    # unbound first (wherever the scheduler put this thread), then from the GPU's own NUMA node
    sys.path.insert(0, ROOT)
    from bench import gpu_local_cpus
    xu = torch.rand(n, dtype=torch.float64).pin_memory()
    du = torch.empty(n, dtype=torch.float64, device="cuda")

    def h2d_unbound():
        du.copy_(xu, non_blocking=True)
        torch.cuda.synchronize()
    out["h2d_unbound_us"] = wall(h2d_unbound, args.reps)
    near = gpu_local_cpus(0)
    out["gpu_local_cpus"] = len(near) if near else 0
    out["all_cpus"] = len(os.sched_getaffinity(0))
    if near:
        os.sched_setaffinity(0, near)
    xh = torch.rand(n, dtype=torch.float64).pin_memory()
    yh = torch.zeros(n, dtype=torch.float64).pin_memory()
    xd = torch.empty(n, dtype=torch.float64, device="cuda")
    yd = torch.ones(n, dtype=torch.float64, device="cuda")
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()

    def h2d():
        xd.copy_(xh, non_blocking=True)
        torch.cuda.synchronize()

    def d2h():
        yh.copy_(yd, non_blocking=True)
        torch.cuda.synchronize()

    def both():
        with torch.cuda.stream(s1):
            xd.copy_(xh, non_blocking=True)
        with torch.cuda.stream(s2):
            yh.copy_(yd, non_blocking=True)
        torch.cuda.synchronize()

    out["h2d_us"] = wall(h2d, args.reps)
    out["d2h_us"] = wall(d2h, args.reps)
    out["h2d_and_d2h_concurrent_us"] = wall(both, args.reps)
    h = Handle(0)
    modes = [int(v) for v in os.environ.get("E2E_MODES", "1,0,2").split(",")]
    chunk_list = [int(v) for v in os.environ.get("E2E_CHUNKS", "1,2,4,6,8,12,16,32").split(",")]
    for direct in modes:
        h.set_option("HOST_Y_DIRECT", direct if direct < 2 else 0)
        h.set_option("HOST_GRAPH", 2 if direct < 2 else 0)
        for chunks in chunk_list:
            h.set_option("HOST_CHUNKS", chunks)
            mat = h.matrix_upload(a, "d")
            xn, yn = xh.numpy(), yh.numpy()
            out["spmv_host_%s_chunks_%d_us" % ({1: "ydirect", 0: "ycopy", 2: "ycopy_graph"}[direct], chunks)] = wall(lambda: h.matrix_spmv_host(mat, xn, yn), args.reps)
            h.matrix_free(mat)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
