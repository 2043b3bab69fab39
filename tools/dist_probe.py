#!/usr/bin/env python
"""Strong-scaling probe at small rank counts: every rank owns `planes` planes of a 134 x 134 x (planes * N) 27-point grid
(planes = 17 is a rank's share of BASELINE configs[4] on 8 GPUs), timed under option variations.
    torchrun --nproc-per-node 2 tools/dist_probe.py [--planes 17] [--steps 400]"""
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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--planes", type=int, default=17)
    ap.add_argument("--steps", type=int, default=400)
    ap.add_argument("--combos", default="")
    ap.add_argument("--modes", default="full,plain,fuse_only,wait_only")
    ap.add_argument("--proto", default="auto", choices=["auto", "pull", "packets", "packets-fused", "flags"])
    ap.add_argument("--packets", type=int, default=1, help="1: {value, step} packet exchange (rsg_dist_enable_packets); 0: arrival flags")
    args = ap.parse_args()
    world, rank, lr = int(os.environ["WORLD_SIZE"]), int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(lr)
    dev = torch.device("cuda", lr)
    os.environ.pop("NCCL_DEBUG", None)
    dist.init_process_group("nccl", device_id=dev)
    G = 134
    nz = args.planes * world
    a = host.gen_stencil27(G, G, nz, rank * args.planes, (rank + 1) * args.planes, seed=2)
    starts = np.arange(world + 1, dtype=np.int64) * a.m
    xg = orc.dlarnv(G * G * nz) - 0.5
    combos = [c for c in args.combos.split(";") if c] or ["", "SPMV_PDL=1", "SPMV_PDL=2", "SPMV_VARIANT=0,SPMV_WIN_BYTES=36864", "SPMV_PDL=1,SPMV_VARIANT=0,SPMV_WIN_BYTES=36864",
                                                           "SPMV_PDL=2,SPMV_VARIANT=0,SPMV_WIN_BYTES=36864", "SPMV_PDL=2,SPMV_VARIANT=1,SPMV_WIN_BYTES=22528",
                                                           "SPMV_PDL=2,SPMV_VARIANT=2,SPMV_WIN_BYTES=75776"]
    for spec in combos:
        h = Handle(lr)
        hp = Handle(lr)
        opts = dict(kv.split("=") for kv in spec.split(",") if kv)
        for k, v in opts.items():
            h.set_option(k, int(v))
        ops = []
        for _ in range(2):
            op = RowBlockSpmv(a, starts, "p2p", lambda c: DeviceCSR(c, "d", dev), lambda A, x, y: h.spmv(A, x, y), dev, torch.float64,
                              push=(hp.halo_push, h.spmv_halo_wait, h.spmv_halo_fuse), proto=args.proto)
            op.set_x(torch.from_numpy(np.ascontiguousarray(xg[starts[rank]:starts[rank + 1]])).to(dev))
            ops.append(op)
        for i in range(10):
            ops[i % 2].step()
        torch.cuda.synchronize()
        res = {}
        scratch = torch.zeros(ops[0].plan.n_ext + 64, dtype=torch.float64, device=dev)
        lflags = torch.zeros(64, dtype=torch.int32, device=dev)
        local_slots = h.ptr_array([scratch.data_ptr() if q != rank else 0 for q in range(world)])
        local_flags = h.ptr_array([lflags[q:].data_ptr() if q != rank else 0 for q in range(world)])
        for mode in args.modes.split(","):
            def one(i, mode=mode):
                op = ops[i % 2]
                if mode == "full":
                    op.step()
                    return
                h.set_option("SPMV_HALO_START", op.plan.halo_base)
                if mode == "fuse_only":  # push + signal with an epoch the peers have long passed: nobody waits on it
                    h.spmv_halo_fuse(op.send_idx32, op.slots[0], local_flags, 1)  # (signals go to a dummy: the real flags count)
                elif mode == "fuse_local":  # the same push, but into a buffer on THIS device: is the NVLink store the slow part?
                    h.spmv_halo_fuse(op.send_idx32, local_slots, local_flags, 1)
                elif mode == "wait_only":  # flags already at a later epoch: the polls succeed at once
                    h.spmv_halo_wait(op.flag_views[0], op.n_wait, 1)
                h.spmv(op.A, op.images[0], op.y)
                h.set_option("SPMV_HALO_START", 0)
            for i in range(6):
                one(i)
            best = None
            for _ in range(3):
                dist.barrier()
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for i in range(args.steps):
                    one(i)
                e1.record()
                torch.cuda.synchronize()
                t = torch.tensor([e0.elapsed_time(e1) / args.steps * 1e3], dtype=torch.float64, device=dev)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                best = float(t.item()) if best is None else min(best, float(t.item()))
            res[mode] = round(best, 2)
            torch.cuda.synchronize()
            dist.barrier()
        best = res.get("full", 0.0)
        for i in range(4):  # leave the operators in a consistent state for the parity check
            ops[i % 2].step()
        h.synchronize()
        ref = orc.spmv(a, xg, "d", threads=8)
        ok = bool(np.array_equal(ops[0].y.cpu().numpy(), ref))
        if rank == 0:
            print(json.dumps({"world": world, "planes": args.planes, "opts": spec, "us_per_step": round(best, 2), "rows": a.m, "stored": a.stored,
                              "tiles": h.info("SPMV_ROW_BLOCKS"), "algo": h.info("SPMV_ALGO_USED"), "bit_exact": ok, "proto": ops[0].describe()["protocol"], "modes": res}), flush=True)
        del ops
        dist.barrier()
    dist.barrier()
    torch.cuda.synchronize()
    sys.stdout.flush()
    os._exit(0)


if __name__ == "__main__":
    main()
