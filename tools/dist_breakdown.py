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
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


for mode in ("p2p", "halo"):
    ops = []
    for r in range(2):
        op = RowBlockSpmv(a, row_starts, mode, lambda c: DeviceCSR(c, "d", dev), lambda A, x, y: h.spmv(A, x, y), dev,
                          torch.float64, push=h.halo_push)
        op.set_x(torch.from_numpy(xs).to(dev))
        h.spmv_analysis(op.A)
        ops.append(op)
    t_spmv = timed(lambda i: h.spmv(ops[i % 2].A, ops[i % 2].x_ext, ops[i % 2].y))
    t_exch = timed(lambda i: ops[i % 2].exchange())
    t_both = timed(lambda i: ops[i % 2].step())
    extra = ""
    if mode == "p2p":
        t_push = timed(lambda i: h.halo_push(ops[0].send_idx32, ops[0].x_ext, ops[0].slots[0]))
        t_bar = timed(lambda i: ops[0].symm.barrier(channel=0))
        extra = " push %.1f barrier %.1f" % (t_push, t_bar)
    if rank == 0:
        print("%s: spmv %.1f us  exchange %.1f us  step %.1f us%s" % (mode, t_spmv, t_exch, t_both, extra), flush=True)
dist.barrier()
torch.cuda.synchronize()
os._exit(0)
