"""World-size-2 (and 3) gloo runs of the row-block partitioned SpMV's host logic on CPU: partition,
halo discovery, column relabelling and the exchange step.  The local product is done by the oracle
here (the CUDA kernel is exercised by the -m gpu tests); the distributed result must equal the
single-process oracle product bit for bit, because relabelling keeps each row's summation order."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, mode, kind, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import orc
    from respasol_b200 import host
    from respasol_b200.dist import RowBlockSpmv
    if kind == "stencil":
        full = host.gen_stencil27(7, 6, 12, seed=2)
    else:
        full = host.gen_powerlaw(3000, 2.2, 24, 2, 700, 4)
    starts = orc.row_partition(full, world, True)
    local = full.rows(int(starts[rank]), int(starts[rank + 1]))
    x = orc.dlarnv(full.n) - 0.5

    def spmv(A, xe, y):
        y.copy_(torch.from_numpy(orc.spmv(A, xe.numpy(), "d")))

    # the stencil runs use the overlap split (interior rows first, boundary pieces after the exchange)
    view = (lambda A, a, b: A.rows(a, b)) if kind == "stencil" else None
    op = RowBlockSpmv(local, starts, mode, lambda c: c, spmv, torch.device("cpu"), torch.float64, view=view)
    if kind == "stencil" and mode == "halo" and world == 2:
        assert op.split is not None and sum(b - a for _, a, b in op.split) == local.m
    op.set_x(torch.from_numpy(x[starts[rank]:starts[rank + 1]].copy()))
    y = op.step().numpy()
    ref = orc.spmv(full, x, "d")[starts[rank]:starts[rank + 1]]
    ok = np.array_equal(y, ref)
    # second product with a changed x slice: the exchange must carry the new values
    x2 = x * 1.5 + 0.25
    op.set_x(torch.from_numpy(x2[starts[rank]:starts[rank + 1]].copy()))
    y2 = op.step().numpy()
    ok = ok and np.array_equal(y2, orc.spmv(full, x2, "d")[starts[rank]:starts[rank + 1]])
    if mode == "halo" and kind == "stencil":  # only the neighbouring planes (cuts need not be plane-aligned)
        ok = ok and op.plan.width <= 2 * (7 * 6 + 7 + 1)
    res = torch.tensor([1 if ok else 0])
    dist.all_reduce(res, op=dist.ReduceOp.MIN)
    if rank == 0:
        out.put(int(res.item()))
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
@pytest.mark.parametrize("mode", ["halo", "allgather"])
@pytest.mark.parametrize("kind", ["stencil", "powerlaw"])
def test_rowblock_spmv_gloo(world, mode, kind):
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, mode, kind, out)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert out.get(timeout=5) == 1


def test_halo_plan_single_process():
    from respasol_b200 import host
    from respasol_b200.dist import HaloPlan
    full = host.gen_stencil27(5, 5, 9, seed=0)
    starts = np.array([0, 75, 150, 225])
    plans = [HaloPlan(full.rows(starts[r], starts[r + 1]), starts, r) for r in range(3)]
    needs = [p.need for p in plans]
    for p in plans:
        p.finish(needs)
    # middle rank needs one plane from each neighbour; end ranks need one plane
    assert [len(n) for n in needs] == [25, 50, 25]
    assert plans[1].width == 50 and plans[0].remapped.n == 96 + 3 * 50 and plans[0].halo_base == 96
    # relabelled columns decode back to the original global columns
    for r, p in enumerate(plans):
        loc = full.rows(starts[r], starts[r + 1])
        dec = np.empty(loc.stored, dtype=np.int64)
        c = p.remapped.colidx.astype(np.int64)
        own = c < p.n_local
        assert not np.any((c >= p.n_local) & (c < p.halo_base))  # the alignment gap is never referenced
        dec[own] = c[own] + starts[r]
        o = (c[~own] - p.halo_base) // p.width
        pos = (c[~own] - p.halo_base) % p.width
        dec[~own] = np.array([p.send_idx[oo][pp] + starts[oo] for oo, pp in zip(o, pos)])
        assert np.array_equal(dec, loc.colidx)


def test_p2p_slot_offsets_and_readers():
    """Host logic of the peer-memory exchange: where this rank's halo slot sits inside every peer's
    symmetric buffer, and which peers read from this rank at all."""
    from respasol_b200 import host
    from respasol_b200.dist import HaloPlan, p2p_slot_offsets
    full = host.gen_stencil27(5, 5, 9, seed=0)
    starts = np.array([0, 75, 175, 225])
    plans = [HaloPlan(full.rows(starts[r], starts[r + 1]), starts, r) for r in range(3)]
    needs = [p.need for p in plans]
    for p in plans:
        p.finish(needs)
    assert [p.readers for p in plans] == [[1], [0, 2], [1]]
    assert all(p.neighbours == [[1], [0, 2], [1]] for p in plans)
    w = plans[0].width
    stride = 512
    offs = p2p_slot_offsets(starts, w, 1, stride)
    # image 0: behind peer q's own entries, slot number = the sender's rank; image 1 one stride further
    assert offs[0] == [96 + w, 128 + w, 64 + w] and offs[1] == [o + stride for o in offs[0]]
    # the slot is exactly where the relabelled columns of the reader point
    p0 = plans[0]
    c = p0.remapped.colidx.astype(np.int64)
    halo_cols = np.unique(c[c >= p0.halo_base])
    assert halo_cols.min() >= offs[0][0] and halo_cols.max() < offs[0][0] + w


def test_exchange_pattern_symmetry_decides_the_packet_protocol():
    """The packet exchange is only safe when every rank reads from every rank it sends to (that mutual dependency is what keeps a
    rank from running two steps ahead of a reader): HaloPlan.symmetric, identical on all ranks."""
    from respasol_b200 import host
    from respasol_b200.dist import PACKETS_BELOW_ENTRIES, HaloPlan
    full = host.gen_stencil27(5, 5, 9, seed=0)
    starts = np.array([0, 75, 150, 225])
    plans = [HaloPlan(full.rows(starts[r], starts[r + 1]), starts, r) for r in range(3)]
    needs = [p.need for p in plans]
    assert all(p.finish(needs).symmetric for p in plans)
    # an upwind pattern: rank r reads only from rank r - 1 (rank 0 feeds, reads nothing)
    n, band = 300, 40
    rp = np.zeros(n + 1, dtype=np.int32)
    rp[1:] = np.cumsum(np.where(np.arange(n) < band, 1, 2))
    ci = np.concatenate([[r] if r < band else [r - band, r] for r in range(n)]).astype(np.int32)
    up = host.CSR(n, n, int(rp[-1]), rp, ci, np.ones(int(rp[-1])), 0)
    starts = np.array([0, 100, 200, 300])
    plans = [HaloPlan(up.rows(starts[r], starts[r + 1]), starts, r) for r in range(3)]
    needs = [p.need for p in plans]
    assert not any(p.finish(needs).symmetric for p in plans)
    assert [p.readers for p in plans] == [[1], [2], []]
    assert PACKETS_BELOW_ENTRIES == 12_000_000  # the same threshold as kDistPdlMinEntries in csrc/dist.cu
    src = open(__file__.replace("tests/test_dist_cpu.py", "respasol_b200/csrc/dist.cu")).read()
    assert "kDistPdlMinEntries = 12000000" in src


def test_pull_lists_fetch_exactly_what_the_relabelled_columns_read():
    """HaloPlan.pull_lists (the pull exchange's setup): filling the halo of x_ext with owner_x[idx] at `col` for every listed entry
    gives the same product as the global one -- on a stencil (contiguous boundary planes) and on a scattered pattern."""
    import orc
    from respasol_b200 import host
    from respasol_b200.dist import HaloPlan
    for full in (host.gen_stencil27(6, 5, 12, seed=1), host.gen_powerlaw(900, 2.2, 16, 2, 300, 4)):
        P = 3
        starts = np.array([0, full.m // 3 & ~3, 2 * full.m // 3 & ~3, full.m])
        plans = [HaloPlan(full.rows(starts[r], starts[r + 1]), starts, r) for r in range(P)]
        needs = [p.need for p in plans]
        x = orc.dlarnv(full.n) - 0.5
        ref = orc.spmv(full, x, "d")
        for r, p in enumerate(plans):
            p.finish(needs)
            col, idx, lo, hi = p.pull_lists()
            assert len(col) == len(idx) == p.recv_entries and lo[0] == 0 and hi[-1] == len(col)
            x_ext = np.zeros(p.n_ext)
            x_ext[:p.n_local] = x[starts[r]:starts[r + 1]]
            for o in range(P):
                assert o != r or hi[o] == lo[o]
                sl = slice(lo[o], hi[o])
                assert np.all(col[sl] >= p.halo_base + o * p.width) and np.all(col[sl] < p.halo_base + (o + 1) * p.width)
                x_ext[col[sl]] = x[starts[o] + idx[sl]]
            assert np.array_equal(orc.spmv(p.remapped, x_ext, "d"), ref[starts[r]:starts[r + 1]]), r
