#!/usr/bin/env python
"""Where a distributed SpMV step's time goes (run under torchrun, N >= 2):
local SpMV alone, exchange alone, both -- CUDA-event time per step, max over ranks.
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/dist_breakdown.py"""
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

world, rank, lr = int(os.environ["WORLD_SIZE"]), int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
h = Handle(lr)
G = 100
a = host.gen_stencil27(G, G, G * world, rank * G, (rank + 1) * G, seed=2)
xs = orc.dlarnv(G * G * G * world)[rank * a.m:(rank + 1) * a.m] - 0.5
row_starts = np.arange(world + 1, dtype=np.int64) * a.m


def timed(fn, K=500):
    for i in range(10):
        fn(i)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    dist.barrier()
    torch.cuda.synchronize()
    e0.record()
    for i in range(K):
        fn(i)
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / K * 1e3], device=dev)
    every = [torch.zeros_like(t) for _ in range(world)]
    dist.all_gather(every, t)
    return [round(float(v.item()), 1) for v in every]


hp = Handle(lr)
side = torch.cuda.Stream(device=dev)
with torch.cuda.stream(side):
    hp.use_current_stream()
for mode in sys.argv[1:] or ("p2p", "p2p-side", "halo"):
    ops = []
    for r in range(2):
        op = RowBlockSpmv(a, row_starts, "p2p" if mode.startswith("p2p") else mode, lambda c: DeviceCSR(c, "d", dev), lambda A, x, y: h.spmv(A, x, y), dev,
                          torch.float64, push=(hp.halo_push, h.spmv_halo_wait, h.spmv_halo_fuse) if mode == "p2p" else (hp.halo_push, h.spmv_halo_wait), side=side)
        op.set_x(torch.from_numpy(xs).to(dev))
        if op.p2p is not None:
            h.set_option("SPMV_HALO_START", op.plan.halo_base)
        h.spmv_analysis(op.A)
        ops.append(op)
    t_spmv = timed(lambda i: h.spmv(ops[i % 2].A, ops[i % 2].x_ext, ops[i % 2].y))
    t_exch = timed(lambda i: ops[i % 2].exchange()) if not mode.startswith("p2p") else "(runs beside the product)"
    t_both = timed(lambda i: ops[i % 2].step())
    h.set_option("SPMV_HALO_START", 0)
    if rank == 0:
        print("%s, us per call on every rank:\n  spmv alone %s\n  exchange alone %s\n  step %s" % (mode, t_spmv, t_exch, t_both), flush=True)
dist.barrier()
torch.cuda.synchronize()
os._exit(0)
