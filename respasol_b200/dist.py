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
  mode "p2p":       same relabelling as "halo", but every rank PUSHES its boundary entries straight into the
                    peers' halo buffers (torch symmetric memory = P2P-mapped allocations over NVLink; the
                    library's rsg_halo_push kernel does the stores), followed by one cross-GPU barrier.
                    x lives in two such buffers used in turn, so a rank one step ahead never overwrites
                    a halo a slower peer is still reading.
Overlap (`view` given): the rows are cut into [0,r0) | [r0,r1) | [r1,m) where [r0,r1) is the longest run
of rows that reference no other rank's entries (for a slab of a stencil: everything but the two outer
planes).  The exchange runs on a side stream while the interior rows are multiplied; the two boundary
pieces follow once the halo has landed.  Row order, and the order inside each row, are unchanged.
ILU0 / SpTRSV do not shard (dependency chains cross any row cut): replicas only.
"""
import numpy as np
import torch
import torch.distributed as dist

from .host import CSR


HALO_ALIGN = 32  # elements: the halo region starts on its own 128/256-byte line, so a cache line never
#                  mixes entries this GPU owns with entries a peer writes while the product is running


PACKETS_BELOW_ENTRIES = 12_000_000  # per-rank entries below which the packet exchange is the default (kDistPdlMinEntries of csrc/dist.cu)


def halo_base(n_local):
    return (int(n_local) + HALO_ALIGN - 1) // HALO_ALIGN * HALO_ALIGN


def p2p_slot_offsets(row_starts, width, rank, stride):
    """Element offset, inside peer q's symmetric buffer (two x_ext images of `stride` elements), of the
    halo slot that belongs to `rank`: image p of peer q holds [its n_local entries, padded | P slots of `width`]."""
    rs = np.asarray(row_starts, dtype=np.int64)
    n_local = np.diff(rs)
    return [[int(p * stride + halo_base(n_local[q]) + rank * width) for q in range(len(n_local))] for p in range(2)]


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
        self.halo_base = halo_base(self.n_local)
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
            new[off] = self.halo_base + own * self.width + pos
        self.n_ext = self.halo_base + P * self.width
        a = self.local
        self.remapped = CSR(a.m, self.n_ext, a.stored, (a.rowptr - a.base).astype(np.int32), new.astype(np.int32),
                            a.values, 0)
        self.my_send = send[self.rank]
        lo, hi = rs[self.rank], rs[self.rank + 1]  # peers that read at least one of this rank's entries
        self.readers = [q for q, nb in enumerate(needs_by_rank) if q != self.rank and np.any((nb >= lo) & (nb < hi))]
        # peer q reads only a part of my send list: the bounding range [lo_q, hi_q) of its entries' positions in it
        # (a slab of a stencil: the lower neighbour reads the first plane, the upper one the last)
        self.send_range = []
        for q, nb in enumerate(needs_by_rank):
            mine = nb[(nb >= lo) & (nb < hi)] - lo if q != self.rank else np.zeros(0, np.int64)
            if len(mine):
                pos = np.searchsorted(self.my_send, mine)
                self.send_range.append((int(pos.min()), int(pos.max()) + 1))
            else:
                self.send_range.append((0, 0))
        self.recv_entries = int(len(self.need))  # halo entries this rank receives per step
        # neighbours[q] (identical on every rank): the ranks q exchanges entries with, in either direction.
        # A rank signals and awaits exactly its neighbours -- enough to keep it within one step of each of them.
        reads = [[o for o in range(P) if o != q and np.any((nb >= rs[o]) & (nb < rs[o + 1]))] for q, nb in enumerate(needs_by_rank)]
        self.neighbours = [sorted(set(reads[q]) | {o for o in range(P) if q in reads[o]}) for q in range(P)]
        # every rank reads from every rank it sends to (and reads from somebody): what the packet exchange needs
        self.symmetric = all(len(reads[q]) > 0 and all(q in reads[o] for o in reads[q]) for q in range(P))
        self.interior = self._interior_run(a, off)
        return self

    def pull_lists(self):
        """The pull exchange's view of the halo (rsg_dist_enable_pull): for every entry this rank reads from another rank, grouped
        by owner -- the halo column it lands in and its index in the owner's x; owner o's entries are [lo[o], hi[o])."""
        rs = self.row_starts
        cols, idxs, lo, hi = [], [], [], []
        n = 0
        for o in range(self.P):
            lo.append(n)
            if o != self.rank:
                mine = self.need[(self.need >= rs[o]) & (self.need < rs[o + 1])] - rs[o]
                pos = np.searchsorted(self.send_idx[o], mine)
                cols.append(self.halo_base + o * self.width + pos)
                idxs.append(mine)
                n += len(mine)
            hi.append(n)
        cat = (lambda v: np.concatenate(v).astype(np.int32)) if cols else (lambda v: np.zeros(0, np.int32))
        return cat(cols), cat(idxs), lo, hi

    @staticmethod
    def _interior_run(a, off):
        """[r0, r1): the longest run of consecutive rows without an off-block column."""
        rp = (a.rowptr - a.base).astype(np.int64)
        per_row = np.add.reduceat(off.astype(np.int64), np.minimum(rp[:-1], max(len(off) - 1, 0))) if len(off) else np.zeros(a.m, np.int64)
        per_row = np.where(rp[1:] > rp[:-1], per_row, 0)  # reduceat on an empty row picks up a neighbour's entry
        dirty = np.flatnonzero(per_row > 0)
        if len(dirty) == 0:
            return 0, a.m
        edges = np.concatenate(([-1], dirty, [a.m]))
        gaps = np.diff(edges) - 1
        g = int(np.argmax(gaps))
        return int(edges[g] + 1), int(edges[g + 1])


class RowBlockSpmv:
    """y_local = A_local * x, x distributed by rows.  `upload(csr)` and `spmv(A, x_ext, y)` are injected so
    the same exchange logic runs on the GPU (libreaspasol_gpu through Handle) and in the gloo CPU tests."""

    def __init__(self, local: CSR, row_starts, mode, upload, spmv, device, dtype, group=None, view=None, push=None, side=None,
                 packets=False, proto=None):
        self.group = group
        # exchange protocol of the rsg_dist mode: "packets" ({value, step} packets pushed by a small kernel on a side stream and
        # unpacked by the halo tiles; symmetric patterns on the window kernel, otherwise the flags stay), "packets-fused" (the
        # product's first CTAs push them), "pull" (the product's first CTAs fetch the halo), "flags" (push + fence + arrival flags)
        self.proto = proto if proto is not None else "auto"
        self.use_packets = self.proto in ("packets", "packets-fused")
        self.split = None
        self.p2p = None
        self.side = side  # p2p: the stream the push kernels run on (the caller binds its push handle to it)
        self.chain = mode == "p2p-chain"
        if mode in ("p2p", "p2p-chain"):
            mode, self.p2p = "halo", push  # (rsg_halo_push, rsg_spmv_halo_wait or None) of the library's handle
            if push[1] is not None:
                view = None  # the boundary tiles wait inside the kernel: nothing to split
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
            r0, r1 = plan.interior
            if view is not None and self.P > 1 and r1 - r0 >= local.m // 2:
                # (piece, first row, end row): interior first, then the pieces that wait for the halo
                self.split = [(view(self.A, a, b), a, b) for a, b in ((r0, r1), (0, r0), (r1, local.m)) if b > a]
                self.interior_rows = r1 - r0
                if torch.device(device).type == "cuda":
                    if self.side is None:
                        self.side = torch.cuda.Stream(device=device)
                    self.landed = torch.cuda.Event()
            if self.p2p is not None and self.P > 1:
                self._setup_p2p(plan, dtype, device)
            else:
                self.p2p = None
                self.x_ext = torch.zeros(plan.n_ext, dtype=dtype, device=device)
            self.send_idx = torch.from_numpy(plan.my_send).to(device)
            self.pack = torch.zeros(plan.width, dtype=dtype, device=device)
            self.halo = self.x_ext[plan.halo_base:]
            # what this rank RECEIVES per step (the entries it reads from other ranks), not the size of the halo buffer
            self.exchange_bytes = plan.recv_entries * self.x_ext.element_size()
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

    def _setup_p2p(self, plan, dtype, device):
        import torch.distributed._symmetric_memory as symm_mem
        group = self.group if self.group is not None else dist.group.WORLD
        n_ext_all = [None] * self.P
        dist.all_gather_object(n_ext_all, int(plan.n_ext), group=self.group)
        stride = (max(n_ext_all) + 31) // 32 * 32
        nimg = 3 if self.chain else 2
        buf = symm_mem.empty(nimg * stride, dtype=dtype, device=device)
        buf.zero_()
        self.symm = symm_mem.rendezvous(buf, group)
        self.symm_buf = buf
        self.images = [buf[p * stride:p * stride + plan.n_ext] for p in range(nimg)]
        self.x_ext = self.images[0]
        esz = buf.element_size()
        offs = p2p_slot_offsets(self.row_starts, plan.width, self.rank, stride)
        ptrs = list(self.symm.buffer_ptrs)
        self.slots = [[(ptrs[q] + offs[p][q] * esz) if q in plan.readers else 0 for q in range(self.P)] for p in range(2)]
        # arrival flags: image p, one int per NEIGHBOUR (slot = the sender's position in the receiver's list).
        # A rank signals all its neighbours, data or not, and awaits all of them, so it can never run more
        # than one step ahead of a rank it exchanges entries with -- the two images are enough
        fl = symm_mem.empty(nimg * 32 + 32, dtype=torch.int32, device=device)  # (+32: pull exchange: ready counter, 2 local flags)
        fl.zero_()
        self.symm_flags = symm_mem.rendezvous(fl, group)
        self.flags = fl
        fptrs = list(self.symm_flags.buffer_ptrs)
        nb = plan.neighbours
        self.n_wait = len(nb[self.rank])
        assert self.n_wait <= 16, "more neighbours than RSG_MAX_PEERS"
        self.flag_slots = [[(fptrs[q] + 4 * (32 * p + nb[q].index(self.rank))) if q in nb[self.rank] else 0
                            for q in range(self.P)] for p in range(2)]
        self.epoch = 0
        self.parity = 0
        if self.side is None:
            self.side = torch.cuda.Stream(device=device)
        to_c = getattr(self.p2p[0], "__self__", None)
        if to_c is not None and hasattr(to_c, "ptr_array"):  # pre-built argument arrays: nothing is rebuilt per step
            self.slots = [to_c.ptr_array(v) for v in self.slots]
            self.flag_slots = [to_c.ptr_array(v) for v in self.flag_slots]
        self.flag_views = [self.flags[32 * p:] for p in range(2)]
        self.send_idx32 = torch.from_numpy(plan.my_send.astype(np.int32)).to(device)
        self.cdist = None
        self.cchain = None
        if self.chain:
            # chained products: three images; the library's rsg_dist_create_chain binds everything, a step is one call
            hnd = self.p2p[2].__self__
            offs3 = [[int(p * stride + halo_base(np.diff(self.row_starts)[q]) + self.rank * plan.width) for q in range(self.P)] for p in range(3)]
            raw_slots = [[(ptrs[q] + offs3[p][q] * esz) if q in plan.readers else 0 for q in range(self.P)] for p in range(3)]
            raw_flags = [[(fptrs[q] + 4 * (32 * p + nb[q].index(self.rank))) if q in nb[self.rank] else 0
                          for q in range(self.P)] for p in range(3)]
            self.flag_views = [self.flags[32 * p:] for p in range(3)]
            self.send_idx32 = torch.from_numpy(plan.my_send.astype(np.int32)).to(device)
            self.cchain = hnd.dist_create_chain(self.A, [im.data_ptr() for im in self.images], plan.halo_base, self.send_idx32,
                                                raw_slots, raw_flags, [r[0] for r in plan.send_range], [r[1] for r in plan.send_range],
                                                [v.data_ptr() for v in self.flag_views], self.n_wait)
            self.chain_hnd = hnd
            self.chain_steps = 0
            torch.cuda.synchronize(device)
            self.symm.barrier(channel=0)
            return
        owner = getattr(self.p2p[0], "__self__", None)
        if len(self.p2p) > 2 and owner is not None and hasattr(self.p2p[2].__self__, "dist_create"):
            # fused mode: everything a step needs is bound once inside the library (rsg_dist_create); a step is ONE call
            hnd = self.p2p[2].__self__
            raw_slots = [[(ptrs[q] + offs[p][q] * esz) if q in plan.readers else 0 for q in range(self.P)] for p in range(2)]
            raw_flags = [[(fptrs[q] + 4 * (32 * p + nb[q].index(self.rank))) if q in nb[self.rank] else 0
                          for q in range(self.P)] for p in range(2)]
            self.cdist = hnd.dist_create(self.A, [im.data_ptr() for im in self.images], plan.halo_base, self.send_idx32,
                                         raw_slots, raw_flags, [r[0] for r in plan.send_range], [r[1] for r in plan.send_range],
                                         [v.data_ptr() for v in self.flag_views], self.n_wait)
            self.cdist_step = hnd.dist_step
            if self.proto == "auto":
                # short products (a rank's share of config 5 on 8 GPUs: 14 us) cannot hide the push -> fence -> counter -> flag chain
                # behind their interior tiles: packets.  Longer ones can, and the flags cost their halo tiles nothing: flags.
                # (measured on config 5: 8 GPUs 16.4 us with packets / 20.4 with flags; 4 GPUs 28.9 / 27.0-27.6)
                big = torch.tensor([int(self.A.stored)], dtype=torch.int64, device=device)
                dist.all_reduce(big, op=dist.ReduceOp.MAX, group=self.group)
                self.proto = "packets" if int(big.item()) < PACKETS_BELOW_ENTRIES else "flags"
                self.use_packets = self.proto == "packets"
            self.pull = False
            if self.proto == "pull":
                # pull exchange (rsg_dist_enable_pull): entries this rank reads, grouped by owner
                pcol, pidx, lo, hi = plan.pull_lists()
                self.pull_col = torch.from_numpy(pcol).to(device)
                self.pull_idx = torch.from_numpy(pidx).to(device)
                base = nimg * 32
                peer_x = [[(ptrs[q] + p * stride * esz) if hi[q] > lo[q] else 0 for q in range(self.P)] for p in range(2)]
                peer_ready = [(fptrs[q] + 4 * base) if q in nb[self.rank] else 0 for q in range(self.P)]
                ok = hnd.dist_enable_pull(self.cdist, self.pull_col, self.pull_idx, peer_x, peer_ready, lo, hi,
                                          self.flags[base:].data_ptr(), self.flags[base + 1:].data_ptr())
                agree = torch.tensor([1 if ok else 0], dtype=torch.int32, device=device)
                dist.all_reduce(agree, op=dist.ReduceOp.MIN, group=self.group)
                assert int(agree.item()) == 1 or not ok, "pull exchange enabled on some ranks only"
                self.pull = bool(ok)
            # packet exchange (rsg_dist_enable_packets): {value, step} packets instead of data + fence + counter + flag
            self.packets = False
            if self.use_packets and plan.symmetric:
                per_image = self.P * plan.width
                pk = symm_mem.empty(2 * per_image * 2, dtype=dtype, device=device)  # a packet = 2 elements
                pk.zero_()
                torch.cuda.synchronize(device)  # the zero fill precedes the marks rsg_dist_enable_packets writes on the handle's stream
                self.symm_pk = symm_mem.rendezvous(pk, group)
                self.pk = pk
                pptrs = list(self.symm_pk.buffer_ptrs)
                pb = 2 * esz
                raw_pk = [[(pptrs[q] + pb * (p * per_image + self.rank * plan.width)) if q in plan.readers else 0
                           for q in range(self.P)] for p in range(2)]
                mine = [pk.data_ptr() + pb * p * per_image for p in range(2)]
                ok = hnd.dist_enable_packets(self.cdist, mine, raw_pk, side_stream=self.proto == "packets")
                agree = torch.tensor([1 if ok else 0], dtype=torch.int32, device=device)
                dist.all_reduce(agree, op=dist.ReduceOp.MIN, group=self.group)
                if int(agree.item()) == 1:
                    self.packets = True
                elif ok:  # some rank's block is not a window plan: everybody back to the flags
                    hnd.dist_enable_packets(self.cdist, None, None)
        torch.cuda.synchronize(device)
        self.symm.barrier(channel=0)

    # ---- chained products (mode "p2p-chain")
    def set_x_global(self, x_global):
        """(re)start the chain from the GLOBAL vector (numpy): image 1 gets this rank's slice and the halo entries it reads"""
        assert self.cchain is not None
        dev = self.x_ext.device
        torch.cuda.synchronize(dev)
        self.symm.barrier(channel=0)  # nobody is still pushing into anybody's images
        plan, rs = self.plan, self.row_starts
        img = self.images[1]
        img[:self.n_local].copy_(torch.from_numpy(np.ascontiguousarray(x_global[rs[self.rank]:rs[self.rank + 1]])).to(dev))
        for o in range(self.P):
            if o != self.rank and len(plan.send_idx[o]):
                vals = np.ascontiguousarray(x_global[plan.send_idx[o] + rs[o]])
                img[plan.halo_base + o * plan.width:plan.halo_base + o * plan.width + len(vals)].copy_(torch.from_numpy(vals).to(dev))
        self.flags.zero_()
        self.chain_hnd.dist_chain_reset(self.cchain)
        self.chain_steps = 0
        torch.cuda.synchronize(dev)
        self.symm.barrier(channel=0)

    def step_chain(self, alpha):
        """x <- alpha * A * x: one library call, one launch (boundary rows pushed by the tiles that computed them)"""
        self.chain_hnd.dist_step_chain(self.cchain, alpha)
        self.chain_steps += 1

    def x_local(self):
        """this rank's slice of the chain's current x"""
        return self.images[(self.chain_steps + 1) % 3][:self.n_local]

    def describe(self):
        """what the exchange step of this operator is, for the bench line"""
        if getattr(self, "cchain", None) is not None:
            mode = ("chained products x <- alpha A x (rsg_dist_step_chain): the tiles that own rows a neighbour reads store them straight "
                    "into the neighbour's halo slots of the next x image over NVLink and the last of them signals; the next product's "
                    "halo tiles acquire the flags before they copy x; three x images rotate; ONE launch per step, no separate exchange")
        elif self.p2p is not None:
            how = ("FETCHED by the SpMV launch itself (rsg_dist_step after rsg_dist_enable_pull: the product's first CTAs poll the "
                   "neighbours' step counters over NVLink, load this rank's halo entries from the owners' x and raise a local flag the "
                   "halo tiles wait for; the product contains no remote store)" if getattr(self, "pull", False) else
                   "as {value, step} PACKETS (rsg_dist_step after rsg_dist_enable_packets: one call = a small push kernel on a side stream "
                   "that stores every boundary entry with the step number in one 16-byte store into the readers' packet buffers, and "
                   "the product, whose halo tiles unpack their x window out of the packets, polling the step number inside each "
                   "entry -- no fence, counter or flag, no remote store inside the product)" if getattr(self, "packets", False) and self.proto == "packets" else
                   "FUSED into the SpMV launch as {value, step} PACKETS (the product's first CTAs push them)" if getattr(self, "packets", False) else
                   "FUSED into the SpMV launch (rsg_dist_step: the product's first CTAs push this rank's boundary entries into the "
                   "readers' halo slots and signal, its halo tiles await the neighbours' flags before they copy x)"
                   if len(self.p2p) > 2 else "rsg_halo_push on a side stream, arrival flags awaited inside the product's halo tiles"
                   if self.p2p[1] else "rsg_halo_push + symmetric-memory barrier")
            mode = "x-halo push over NVLink peer memory, " + how
        elif self.mode == "halo":
            mode = "packed x-halo all-gather (NCCL)"
        else:
            mode = "full x all-gather (NCCL)"
        if self.split:
            mode += "; exchange on a side stream overlapped with the %d interior rows, boundary rows follow" % self.interior_rows
        return {"mode": mode, "bytes_received_per_rank_per_step": int(self.exchange_bytes),
                "neighbours": int(getattr(self, "n_wait", 0)), "protocol": "pull" if getattr(self, "pull", False) else "packets" if getattr(self, "packets", False) else "flags"}

    def set_x(self, x_local):
        """x_local: this rank's slice of x (device tensor)."""
        if self.p2p is not None:
            for im in self.images:
                im[:self.n_local].copy_(x_local)
        elif self.mode == "halo":
            self.x_ext[:self.n_local].copy_(x_local)
        else:
            self.pack[:self.n_local].copy_(x_local)

    def exchange(self):
        if self.p2p is not None:
            # push + signal; the product that follows waits for the peers' flags inside its boundary tiles
            self.epoch += 1
            self.x_ext = self.images[self.parity]
            push, arm = self.p2p[0], self.p2p[1]
            fuse = self.p2p[2] if len(self.p2p) > 2 else None
            if fuse is not None:
                # ONE launch: the product's first CTAs push this rank's boundary entries and signal, its boundary
                # tiles wait for the neighbours' signals, everything else multiplies meanwhile
                fuse(self.send_idx32, self.slots[self.parity], self.flag_slots[self.parity], self.epoch)
                arm(self.flag_views[self.parity], self.n_wait, self.epoch)
            elif arm is not None:
                # the push runs beside the product (its own stream): only the boundary tiles wait for a peer
                # (`push` launches on self.side: the caller's push handle is bound to that stream)
                self.side.wait_stream(torch.cuda.current_stream(self.x_ext.device))
                push(self.send_idx32, self.x_ext, self.slots[self.parity], self.flag_slots[self.parity], self.epoch)
                arm(self.flag_views[self.parity], self.n_wait, self.epoch)
                self.join_side = True
            else:
                push(self.send_idx32, self.x_ext, self.slots[self.parity], self.flag_slots[self.parity], self.epoch)
                self.symm.barrier(channel=0)  # no in-kernel wait: every peer's push into this image has landed
            self.parity ^= 1
        elif self.mode == "halo":
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

    def capture(self, rebind_stream):
        """Record one step (pack, all-gather, the SpMV launches, the stream fork/join) into a CUDA graph:
        a step is 5-7 launches of 5-60 us each, and issuing them from Python costs more than they run.
        `rebind_stream()` points the SpMV library's handle at torch's current stream."""
        self.graph = None
        torch.cuda.synchronize(self.x_ext.device)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            rebind_stream()
            self.step()
        rebind_stream()
        self.graph = g
        return g

    def step(self):
        """one distributed product: exchange step, then the local SpMV into self.y"""
        if getattr(self, "graph", None) is not None:
            self.graph.replay()
            return self.y
        if self.split is None:
            if getattr(self, "cdist", None) is not None:  # push + signal + wait + product: one library call, one launch
                self.cdist_step(self.cdist, self.y)
                self.epoch += 1
                self.x_ext = self.images[self.parity]
                self.parity ^= 1
                return self.y
            self.exchange()
            self.spmv_fn(self.A, self.x_ext, self.y)
            if getattr(self, "join_side", False):  # x may be rewritten once the step returns
                torch.cuda.current_stream(self.x_ext.device).wait_stream(self.side)
            return self.y
        cuda = self.x_ext.is_cuda
        if cuda:
            main = torch.cuda.current_stream(self.x_ext.device)
            self.side.wait_stream(main)  # this rank's slice of x is final
            with torch.cuda.stream(self.side):
                self.exchange()
                self.landed.record(self.side)
        else:
            self.exchange()
        for j, (piece, a, b) in enumerate(self.split):
            if j == 1 and cuda:
                main.wait_event(self.landed)
            self.spmv_fn(piece, self.x_ext, self.y[a:b])
        if cuda and len(self.split) == 1:
            main.wait_event(self.landed)
        return self.y
