"""Row-block partitioned SpMV across ranks (one process per GPU, torch.distributed: NCCL over
NVLink/NVSwitch on the box, gloo in the CPU tests).  New capability asked for by BASELINE.json
north_star -- the reference is single-GPU.

Each rank owns a contiguous row block [row_starts[r], row_starts[r+1]) of A together with the
matching slice of x.  One product = one exchange step + one local SpMV:
  mode "allgather": every rank's x slice is all-gathered into a full-length x (padded equal slices)
  mode "halo":      only the x entries some other rank references are packed and all-gathered;
                    off-block column indices are remapped ONCE at setup into the packed halo
                    buffer appended behind the local slice.  Relabelling keeps the order of the
                    entries inside each row, so every row is summed in the same order as on one GPU.
ILU0 / SpTRSV do not shard (dependency chains cross any row cut): replicas only.
"""
import numpy as np
import torch
import torch.distributed as dist

from .host import CSR


def owner_of(cols, row_starts):
    return np.searchsorted(row_starts, cols, side="right") - 1


class HaloPlan:
    """Pure host logic (numpy): which entries each rank must publish, and the relabelled columns."""

    def __init__(self, local: CSR, row_starts, rank, needs_by_rank=None):
        self.rank = rank
        self.P = len(row_starts) - 1
        self.row_starts = np.asarray(row_starts, dtype=np.int64)
        self.lo, self.hi = int(row_starts[rank]), int(row_starts[rank + 1])
        self.n_local = self.hi - self.lo
        cols = local.colidx.astype(np.int64) - local.base
        off = (cols < self.lo) | (cols >= self.hi)
        self.need = np.unique(cols[off])  # global indices this rank reads from other ranks
        self._cols, self._off = cols, off
        self.local = local

    def finish(self, needs_by_rank):
        """needs_by_rank[q] = global indices rank q needs (from anyone).  Deterministic on all ranks."""
        P, rs = self.P, self.row_starts
        send = []
        for o in range(P):  # entries of owner o that anybody needs, ascending, as owner-local indices
            parts = [nb[(nb >= rs[o]) & (nb < rs[o + 1])] for q, nb in enumerate(needs_by_rank) if q != o]
            u = np.unique(np.concatenate(parts)) if parts else np.zeros(0, np.int64)
            send.append((u - rs[o]).astype(np.int64))
        self.send_idx = send
        self.width = max(1, max(len(s) for s in send))  # equal contribution per rank for the all-gather
        cols, off = self._cols, self._off
        new = cols - self.lo
        if off.any():
            oc = cols[off]
            own = owner_of(oc, rs)
            pos = np.empty(len(oc), dtype=np.int64)
            for o in np.unique(own):
                sel = own == o
                pos[sel] = np.searchsorted(send[o], oc[sel] - rs[o])
            new[off] = self.n_local + own * self.width + pos
        self.n_ext = self.n_local + P * self.width
        a = self.local
        self.remapped = CSR(a.m, self.n_ext, a.stored, (a.rowptr - a.base).astype(np.int32), new.astype(np.int32),
                            a.values, 0)
        self.my_send = send[self.rank]
        return self


class RowBlockSpmv:
    """y_local = A_local * x, x distributed by rows.  `upload(csr)` and `spmv(A, x_ext, y)` are injected so
    the same exchange logic runs on the GPU (libreaspasol_gpu through Handle) and in the gloo CPU tests."""

    def __init__(self, local: CSR, row_starts, mode, upload, spmv, device, dtype, group=None):
        self.group = group
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.P = len(row_starts) - 1
        self.mode = mode
        self.spmv_fn = spmv
        self.device = device
        self.row_starts = np.asarray(row_starts, dtype=np.int64)
        self.n_local = int(row_starts[self.rank + 1] - row_starts[self.rank])
        if mode == "halo":
            plan = HaloPlan(local, row_starts, self.rank)
            if self.P > 1:
                needs = [None] * self.P
                dist.all_gather_object(needs, plan.need, group=group)
            else:
                needs = [plan.need]
            plan.finish(needs)
            self.plan = plan
            self.A = upload(plan.remapped)
            self.x_ext = torch.zeros(plan.n_ext, dtype=dtype, device=device)
            self.send_idx = torch.from_numpy(plan.my_send).to(device)
            self.pack = torch.zeros(plan.width, dtype=dtype, device=device)
            self.halo = self.x_ext[plan.n_local:]
            self.exchange_bytes = plan.width * self.P * self.x_ext.element_size()
        elif mode == "allgather":
            self.width = int(np.diff(self.row_starts).max())
            # padded equal slices: global column c of owner o lives at o*width + (c - start[o])
            cols = local.colidx.astype(np.int64) - local.base
            own = owner_of(cols, self.row_starts)
            new = own * self.width + (cols - self.row_starts[own])
            rem = CSR(local.m, self.P * self.width, local.stored, (local.rowptr - local.base).astype(np.int32),
                      new.astype(np.int32), local.values, 0)
            self.A = upload(rem)
            self.x_ext = torch.zeros(self.P * self.width, dtype=dtype, device=device)
            self.pack = torch.zeros(self.width, dtype=dtype, device=device)
            self.exchange_bytes = self.width * self.P * self.x_ext.element_size()
        else:
            raise ValueError(mode)
        self.y = torch.zeros(local.m, dtype=dtype, device=device)

    def set_x(self, x_local):
        """x_local: this rank's slice of x (device tensor)."""
        if self.mode == "halo":
            self.x_ext[:self.n_local].copy_(x_local)
        else:
            self.pack[:self.n_local].copy_(x_local)

    def exchange(self):
        if self.mode == "halo":
            if self.P == 1:
                return
            n = self.send_idx.numel()
            if n:
                torch.index_select(self.x_ext[:self.n_local], 0, self.send_idx, out=self.pack[:n])
            dist.all_gather_into_tensor(self.halo, self.pack, group=self.group)
        else:
            if self.P == 1:
                self.x_ext.copy_(self.pack)
            else:
                dist.all_gather_into_tensor(self.x_ext, self.pack, group=self.group)

    def step(self):
        """one distributed product: exchange step, then the local SpMV into self.y"""
        self.exchange()
        self.spmv_fn(self.A, self.x_ext, self.y)
        return self.y
