#!/usr/bin/env python
"""Per-CTA timeline of the row-block product under the exchange (rsg_debug_trace): where do the microseconds between the plain
product and the exchanged one go?  torchrun --nproc-per-node N tools/dist_trace.py [--planes 17] [--packets 1] [--opts SPMV_PDL=2]
Prints, per rank, for the LAST product of a run of back-to-back steps (times in us relative to the first CTA entry of that
product): when the previous product ended, when the push CTAs ran, and entry / dependency / staged / done percentiles of the
interior tiles and of the tiles that read a peer's entries."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import orc  # noqa: E402
from respasol_b200 import host  # noqa: E402
from respasol_b200.dist import RowBlockSpmv  # noqa: E402
from respasol_b200.gpu import DeviceCSR, Handle  # noqa: E402


def pct(a, qs=(0, 10, 50, 90, 100)):
    return [round(float(np.percentile(a, q)), 2) for q in qs] if len(a) else []


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--planes", type=int, default=17)
    ap.add_argument("--steps", type=int, default=301)
    ap.add_argument("--proto", default="auto", choices=["auto", "pull", "packets", "packets-fused", "flags"])
    ap.add_argument("--packets", type=int, default=1)
    ap.add_argument("--opts", default="SPMV_PDL=2")
    ap.add_argument("--plain", type=int, default=0, help="1: trace the plain product (no exchange) instead")
    args = ap.parse_args()
    world, rank, lr = int(os.environ["WORLD_SIZE"]), int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(lr)
    dev = torch.device("cuda", lr)
    dist.init_process_group("nccl", device_id=dev)
    G = 134
    nz = args.planes * world
    a = host.gen_stencil27(G, G, nz, rank * args.planes, (rank + 1) * args.planes, seed=2)
    starts = np.arange(world + 1, dtype=np.int64) * a.m
    xg = orc.dlarnv(G * G * nz) - 0.5
    h, hp = Handle(lr), Handle(lr)
    for kv in filter(None, args.opts.split(",")):
        k, v = kv.split("=")
        h.set_option(k, int(v))
    ops = []
    for _ in range(2):
        op = RowBlockSpmv(a, starts, "p2p", lambda c: DeviceCSR(c, "d", dev), lambda A, x, y: h.spmv(A, x, y), dev, torch.float64,
                          push=(hp.halo_push, h.spmv_halo_wait, h.spmv_halo_fuse), proto=args.proto)
        op.set_x(torch.from_numpy(np.ascontiguousarray(xg[starts[rank]:starts[rank + 1]])).to(dev))
        ops.append(op)
    cap = 16384
    buf = torch.zeros(2 * 4 * cap, dtype=torch.int64, device=dev)

    evs = [torch.cuda.Event() for _ in range(8)]

    def one(i):
        op = ops[i % 2]
        if args.plain:
            h.spmv(op.A, op.images[0], op.y)
            if args.plain == 2:  # does an event record between two products cost the programmatic overlap?
                evs[i % 8].record()
        else:
            op.step()
    for i in range(20):
        one(i)
    torch.cuda.synchronize()
    dist.barrier()
    assert h.lib.rsg_debug_trace(h.h, buf.data_ptr(), cap) == 0
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):
        one(i)
    e1.record()
    torch.cuda.synchronize()
    assert h.lib.rsg_debug_trace(h.h, None, 0) == 0
    us = e0.elapsed_time(e1) / args.steps * 1e3
    t = buf.cpu().numpy().astype(np.uint64).reshape(2, cap, 4)
    last = (args.steps - 1) & 1  # half used by the last product (parity starts at 0 with the first traced launch)
    cur, prev = t[last], t[last ^ 1]
    halo = (cur[:, 3] >> np.uint64(63)).astype(bool)
    cur = (cur & np.uint64((1 << 63) - 1)).astype(np.int64)
    prev = (prev & np.uint64((1 << 63) - 1)).astype(np.int64)
    used = cur[:, 0] > 0
    n = int(used.sum())
    cur, halo = cur[:n], halo[:n]
    T0 = cur[:, 0].min()
    rel = (cur - T0) / 1e3
    nfuse = 0
    if not args.plain:
        while nfuse < n and cur[nfuse, 2] == cur[nfuse, 3] and nfuse < 1024 and not halo[nfuse]:
            nfuse += 1  # (push CTAs record staged == done)
    tiles = rel[nfuse:]
    th = halo[nfuse:]
    out = {"rank": rank, "world": world, "us_per_step": round(us, 2), "ctas": n, "push_ctas": nfuse, "halo_tiles": int(th.sum()),
           "prev_product_end": round(float((prev[:n, 3].max() - T0) / 1e3), 2), "this_product_end": round(float(rel[:, 3].max()), 2),
           "first_halo_tile_at_cta": int(np.flatnonzero(th)[0]) if th.any() else -1,
           "last_halo_tile_at_cta": int(np.flatnonzero(th)[-1]) if th.any() else -1}
    if nfuse:
        out["push"] = {"entry": pct(rel[:nfuse, 0]), "dependency": pct(rel[:nfuse, 1]), "done": pct(rel[:nfuse, 3])}
    for name, sel in (("interior", ~th), ("halo", th)):
        x = tiles[sel]
        if len(x):
            out[name] = {"entry": pct(x[:, 0]), "dependency": pct(x[:, 1]), "staged": pct(x[:, 2]), "done": pct(x[:, 3]),
                         "staged_minus_dep": pct(x[:, 2] - np.maximum(x[:, 1], x[:, 0])), "life": pct(x[:, 3] - x[:, 0])}
    for r in range(world):
        if r == rank:
            print(json.dumps(out), flush=True)
        dist.barrier()
    torch.cuda.synchronize()
    os._exit(0)


if __name__ == "__main__":
    main()
