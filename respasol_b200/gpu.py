"""Handle: libreaspasol_gpu's C ABI called on torch CUDA tensors.

torch is used for device memory and streams only; every kernel that runs is one of this
repository's (respasol_b200/csrc).  Method names and argument meaning follow the C ABI
(include/reaspasol_gpu.h), which in turn mirrors the cuSPARSE call groups of the reference
harness (GPU/spmv.cu:122-186, GPU/ilu0.cu:196-310)."""
import ctypes as C

import numpy as np
import torch

from ._lib import gpu_lib
from .host import CSR

OPT = dict(SPMV_ALGO=0, SPMV_REDUCE=1, SPMV_LONG_ROW=2, SPMV_VARIANT=3, SPMV_CACHE_HINT=4,
           TRSV_ALGO=5, ILU0_ALGO=6, SPMV_X_PERSIST=7, SPMV_X_HINT=8,
           SPMV_ROWS_PER_THREAD=9, SPMV_CTAS_PER_SM=10, TRSV_BATCH=11, TRSV_GATE=12,
           TRI_CTAS_PER_SM=13, TRSV_BLOCK=14, SPMV_HALO_START=15, HOST_CHUNKS=16, HOST_Y_DIRECT=17, HOST_GRAPH=18, TRANSPOSE_STATIC=19,
           SPMV_WIN_BYTES=20, SPMV_WIN_STREAM_PCT=21, SPMV_WIN_GAP=22, WAIT_SPINS=23, SPMV_PDL=24, SPMV_HALO_PLACE=25, SPMV_HALO_END=26)
ALGO = dict(auto=0, rowblock=1, merge=2, vector=3, scalar=4, stream=5, pipe=6, tile=7, window=8)
INFO = dict(KERNEL_LAUNCHES=0, SPMV_ALGO_USED=1, SPMV_ROW_BLOCKS=2, SPMV_MAX_ROW=3, LEVELS_L=4, LEVELS_U=5,
            SM_COUNT=6, WORKSPACE_BYTES=7)
OK, ZERO_PIVOT_STRUCTURAL, ZERO_PIVOT_NUMERIC = 0, 5, 6

_PREC = {"d": (torch.float64, torch.float64), "s": (torch.float32, torch.float32), "m": (torch.float32, torch.float64)}


class RsgError(RuntimeError):
    def __init__(self, status, where):
        self.status = status
        msg = gpu_lib().rsg_status_string(status).decode()
        super().__init__(f"{where}: {msg} (status {status})")


def _check(status, where, allow=()):
    if status != OK and status not in allow:
        raise RsgError(status, where)
    return status


class DeviceCSR:
    """CSR arrays resident in HBM: int32 rowptr/colidx, values in the storage precision."""
    _next_uid = 0

    def __init__(self, a: CSR, precision="d", device="cuda:0"):
        DeviceCSR._next_uid += 1
        self.uid = DeviceCSR._next_uid
        self.m, self.n, self.base, self.precision = a.m, a.n, a.base, precision
        self.stored = a.stored
        off = int(a.rowptr[0] - a.base)
        if off:  # entries are addressed by rowptr-base: keep a leading gap so indices stay direct
            raise ValueError("row-slab views must be rebased before upload (CSR.rows does this)")
        vt, _ = _PREC[precision]
        self.rowptr = torch.from_numpy(a.rowptr.astype(np.int32)).to(device)
        self.colidx = torch.from_numpy(a.colidx.astype(np.int32)).to(device)
        self.values = torch.from_numpy(a.values).to(device).to(vt)
        self.device = self.rowptr.device

    def view_rows(self, r0, r1):
        """Rows [r0, r1) as a slab view: the row pointers are a slice of the parent's (they keep indexing
        the parent's colidx / values directly, which is what the C ABI's `first = rowptr[0]-base` handles)."""
        v = object.__new__(DeviceCSR)
        DeviceCSR._next_uid += 1
        v.uid = DeviceCSR._next_uid
        v.m, v.n, v.base, v.precision = r1 - r0, self.n, self.base, self.precision
        v.rowptr = self.rowptr[r0:r1 + 1]
        v.colidx, v.values, v.device = self.colidx, self.values, self.device
        v.stored = self.stored  # advisory (the library reads rowptr[m]-rowptr[0])
        return v

    def bytes(self):
        v, x = {"d": (8, 8), "s": (4, 4), "m": (4, 8)}[self.precision]
        return self.stored * (v + 4) + (self.m + 1) * 4 + self.m * x + self.n * x


class Handle:
    def __init__(self, device=0):
        self.lib = gpu_lib()
        self.device = torch.device("cuda", device) if not isinstance(device, torch.device) else device
        h = C.c_void_p()
        _check(self.lib.rsg_create(C.byref(h), self.device.index or 0), "rsg_create")
        self.h = h
        self._owner = {}  # rowptr device address -> uid of the DeviceCSR last seen there
        self.use_current_stream()

    def close(self):
        if getattr(self, "h", None):
            self.lib.rsg_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- plumbing
    def use_current_stream(self):
        s = torch.cuda.current_stream(self.device).cuda_stream
        _check(self.lib.rsg_set_stream(self.h, C.c_void_p(s)), "rsg_set_stream")

    def set_option(self, name, value):
        _check(self.lib.rsg_set_option(self.h, OPT[name], int(value)), "rsg_set_option")

    def info(self, name):
        v = C.c_longlong()
        _check(self.lib.rsg_get_info(self.h, INFO[name], C.byref(v)), "rsg_get_info")
        return v.value

    def synchronize(self):
        _check(self.lib.rsg_synchronize(self.h), "rsg_synchronize")

    @staticmethod
    def _p(t):
        return C.c_void_p(t.data_ptr()) if t is not None else None

    def _claim(self, A):
        """The library caches plans per device address; torch's allocator recycles addresses, so a
        new DeviceCSR appearing at a known rowptr address drops the plans cached for the old one."""
        key = A.rowptr.data_ptr()
        if self._owner.get(key) != A.uid:
            if key in self._owner:
                _check(self.lib.rsg_forget_matrix(self.h, C.c_void_p(key)), "rsg_forget_matrix")
            self._owner[key] = A.uid

    # ---- SpMV
    def spmv_analysis(self, A: DeviceCSR):
        self._claim(A)
        _check(self.lib.rsg_spmv_analysis(self.h, A.m, A.n, A.stored, A.base, self._p(A.rowptr), self._p(A.colidx)),
               "rsg_spmv_analysis")

    def spmv(self, A: DeviceCSR, x, y=None, alpha=1.0, beta=0.0):
        """y = alpha*A*x + beta*y through spmv_csr_d / spmv_csr_s / spmv_csr_mixed."""
        self._claim(A)
        _, xt = _PREC[A.precision]
        assert x.dtype == xt and x.is_cuda and x.numel() >= A.n
        if y is None:
            y = torch.empty(A.m, dtype=xt, device=x.device)
        fn = {"d": self.lib.spmv_csr_d, "s": self.lib.spmv_csr_s, "m": self.lib.spmv_csr_mixed}[A.precision]
        _check(fn(self.h, A.m, A.n, A.stored, A.base, self._p(A.rowptr), self._p(A.colidx), self._p(A.values),
                  alpha, self._p(x), beta, self._p(y)), "spmv_csr_" + A.precision)
        return y

    def spmv_t(self, A: DeviceCSR, x, y=None, alpha=1.0, beta=0.0):
        """y = alpha*A^T*x + beta*y through spmv_csr_t_d / spmv_csr_t_s (x: m entries, y: n entries)."""
        self._claim(A)
        assert A.precision in ("d", "s") and x.dtype == A.values.dtype and x.numel() >= A.m
        if y is None:
            y = torch.empty(A.n, dtype=x.dtype, device=x.device)
        fn = self.lib.spmv_csr_t_d if A.precision == "d" else self.lib.spmv_csr_t_s
        _check(fn(self.h, A.m, A.n, A.stored, A.base, self._p(A.rowptr), self._p(A.colidx), self._p(A.values),
                  alpha, self._p(x), beta, self._p(y)), "spmv_csr_t_" + A.precision)
        return y

    def csr2csc(self, A: DeviceCSR):
        """(colptr, rowidx, values) of A as CSC = CSR of A^T, built on the device."""
        self._claim(A)
        assert A.precision in ("d", "s")
        cnt = int(A.rowptr[-1] - A.rowptr[0])
        colptr = torch.empty(A.n + 1, dtype=torch.int32, device=A.device)
        rowidx = torch.empty(max(cnt, 1), dtype=torch.int32, device=A.device)
        vals = torch.empty(max(cnt, 1), dtype=A.values.dtype, device=A.device)
        fn = self.lib.rsg_csr2csc_d if A.precision == "d" else self.lib.rsg_csr2csc_s
        _check(fn(self.h, A.m, A.n, A.stored, A.base, self._p(A.rowptr), self._p(A.colidx), self._p(A.values),
                  self._p(colptr), self._p(rowidx), self._p(vals)), "rsg_csr2csc")
        return colptr, rowidx[:cnt], vals[:cnt]

    @staticmethod
    def ptr_array(addresses):
        """HOST array of device addresses (0 = NULL) in the form the C ABI takes; build once, pass many times."""
        return (C.c_void_p * len(addresses))(*[C.c_void_p(int(p) if p else None) for p in addresses])

    def halo_push(self, send_idx, x_local, peer_slots, peer_flags=None, epoch=0):
        """x_local[send_idx[j]] -> peer_slots[q][j] for every non-zero device address in peer_slots (the peers'
        P2P-mapped halo buffers, already offset to this rank's slot); then, if peer_flags is given, `epoch`
        is released to every non-zero flag address.  One launch on the handle's stream."""
        n = len(peer_slots)
        slots = peer_slots if isinstance(peer_slots, C.Array) else self.ptr_array(peer_slots)
        flags = None
        if peer_flags is not None:
            assert len(peer_flags) == n and epoch > 0
            flags = peer_flags if isinstance(peer_flags, C.Array) else self.ptr_array(peer_flags)
        _check(self.lib.rsg_halo_push(self.h, x_local.element_size(), int(send_idx.numel()), self._p(send_idx),
                                      self._p(x_local), n, slots, flags, int(epoch)), "rsg_halo_push")

    def spmv_halo_fuse(self, send_idx, peer_slots, peer_flags, epoch):
        """One-shot: the next spmv does the push of halo_push itself (its first CTAs), reading the entries from its own x."""
        slots = peer_slots if isinstance(peer_slots, C.Array) else self.ptr_array(peer_slots)
        flags = peer_flags if isinstance(peer_flags, C.Array) else self.ptr_array(peer_flags)
        _check(self.lib.rsg_spmv_halo_fuse(self.h, int(send_idx.numel()), self._p(send_idx), len(slots), slots, flags,
                                           int(epoch)), "rsg_spmv_halo_fuse")

    def dist_create(self, A, x_images, halo_start, send_idx, peer_slots, peer_flags, peer_lo, peer_hi, wait_flags, nwait):
        """rsg_dist_create: bind one rank's side of the row-block product once.  peer_slots / peer_flags: [image][peer] device
        addresses (0 = NULL); x_images / wait_flags: the two images' device addresses."""
        self._claim(A)
        npeers = len(peer_slots[0])
        assert len(peer_slots) == 2 and len(peer_flags) == 2 and len(peer_flags[0]) == npeers
        slots = self.ptr_array([p for img in peer_slots for p in img])
        flags = self.ptr_array([p for img in peer_flags for p in img])
        lo = (C.c_int * npeers)(*[int(v) for v in peer_lo])
        hi = (C.c_int * npeers)(*[int(v) for v in peer_hi])
        ximg = self.ptr_array(x_images)
        wf = self.ptr_array(wait_flags)
        d = C.c_void_p()
        _check(self.lib.rsg_dist_create(self.h, A.precision.encode(), A.m, A.n, A.stored, A.base, self._p(A.rowptr), self._p(A.colidx),
                                        self._p(A.values), ximg, int(halo_start), int(send_idx.numel()), self._p(send_idx), npeers,
                                        slots, flags, lo, hi, wf, int(nwait), C.byref(d)), "rsg_dist_create")
        return d

    def dist_create_chain(self, A, x_images, halo_start, send_idx, peer_slots, peer_flags, peer_lo, peer_hi, wait_flags, nwait):
        """rsg_dist_create_chain: as dist_create with THREE images (chained products x <- alpha A x)."""
        self._claim(A)
        npeers = len(peer_slots[0])
        assert len(peer_slots) == 3 and len(peer_flags) == 3 and len(x_images) == 3 and len(wait_flags) == 3
        slots = self.ptr_array([p for img in peer_slots for p in img])
        flags = self.ptr_array([p for img in peer_flags for p in img])
        lo = (C.c_int * npeers)(*[int(v) for v in peer_lo])
        hi = (C.c_int * npeers)(*[int(v) for v in peer_hi])
        d = C.c_void_p()
        _check(self.lib.rsg_dist_create_chain(self.h, A.precision.encode(), A.m, A.n, A.stored, A.base, self._p(A.rowptr), self._p(A.colidx),
                                              self._p(A.values), self.ptr_array(x_images), int(halo_start), int(send_idx.numel()),
                                              self._p(send_idx), npeers, slots, flags, lo, hi, self.ptr_array(wait_flags), int(nwait),
                                              C.byref(d)), "rsg_dist_create_chain")
        return d

    def dist_step_chain(self, d, alpha):
        _check(self.lib.rsg_dist_step_chain(d, float(alpha)), "rsg_dist_step_chain")

    def dist_chain_reset(self, d):
        _check(self.lib.rsg_dist_chain_reset(d), "rsg_dist_chain_reset")

    def dist_step(self, d, y):
        _check(self.lib.rsg_dist_step(d, self._p(y)), "rsg_dist_step")

    def dist_enable_packets(self, d, local_packets, peer_packets, side_stream=True):
        """rsg_dist_enable_packets: local_packets = the 2 images' device addresses, peer_packets = [image][peer] (0 = NULL).
        Returns False when the block's plan is not the window kernel's (the flags stay)."""
        if local_packets is None:  # back to the arrival flags
            _check(self.lib.rsg_dist_enable_packets(d, None, None, 0), "rsg_dist_enable_packets")
            return False
        mine = self.ptr_array(local_packets)
        peers = self.ptr_array([p for img in peer_packets for p in img])
        rc = _check(self.lib.rsg_dist_enable_packets(d, mine, peers, 1 if side_stream else 0), "rsg_dist_enable_packets", allow=(8,))
        return rc == OK

    def dist_enable_pull(self, d, pull_col, pull_idx, peer_x, peer_ready, owner_lo, owner_hi, my_ready, local_flags):
        """rsg_dist_enable_pull: peer_x = [image][peer] device addresses (0 = NULL), peer_ready = [peer]; my_ready / local_flags
        device addresses.  Returns False when the block runs a gather kernel (the flags stay)."""
        npeers = len(peer_ready)
        px = self.ptr_array([p for img in peer_x for p in img])
        pr = self.ptr_array(peer_ready)
        lo = (C.c_int * npeers)(*[int(v) for v in owner_lo])
        hi = (C.c_int * npeers)(*[int(v) for v in owner_hi])
        rc = _check(self.lib.rsg_dist_enable_pull(d, int(pull_col.numel()), self._p(pull_col), self._p(pull_idx), npeers, px, pr, lo, hi,
                                                  C.c_void_p(int(my_ready)), C.c_void_p(int(local_flags))),
                    "rsg_dist_enable_pull", allow=(8,))
        return rc == OK

    def dist_destroy(self, d):
        _check(self.lib.rsg_dist_destroy(d), "rsg_dist_destroy")

    def spmv_halo_wait(self, flags, nflags, epoch):
        """One-shot: tiles of the next spmv that read columns >= SPMV_HALO_START wait for flags[i] >= epoch."""
        _check(self.lib.rsg_spmv_halo_wait(self.h, self._p(flags), int(nflags), int(epoch)), "rsg_spmv_halo_wait")

    # ---- ILU0 / SpTRSV
    def ilu0_analysis(self, A: DeviceCSR):
        self._claim(A)
        zp = C.c_int(-1)
        st = self.lib.rsg_ilu0_analysis(self.h, A.m, A.stored, A.base, self._p(A.rowptr), self._p(A.colidx), C.byref(zp))
        _check(st, "rsg_ilu0_analysis", allow=(ZERO_PIVOT_STRUCTURAL,))
        return st, zp.value

    def sptrsv_analysis(self, A: DeviceCSR):
        self._claim(A)
        _check(self.lib.rsg_sptrsv_analysis(self.h, A.m, A.stored, A.base, self._p(A.rowptr), self._p(A.colidx)),
               "rsg_sptrsv_analysis")

    def ilu0_factor(self, A: DeviceCSR, values=None, report_pivot=True):
        """In-place ILU(0) of `values` (default A.values).  Returns (status, zero_pivot_row)."""
        self._claim(A)
        v = A.values if values is None else values
        fn = self.lib.ilu0_factor_d if v.dtype == torch.float64 else self.lib.ilu0_factor_s
        zp = C.c_int(-1)
        st = fn(self.h, A.m, A.stored, A.base, self._p(A.rowptr), self._p(A.colidx), self._p(v),
                C.byref(zp) if report_pivot else None)
        _check(st, "ilu0_factor", allow=(ZERO_PIVOT_STRUCTURAL, ZERO_PIVOT_NUMERIC))
        return st, zp.value

    def sptrsv_lu(self, A: DeviceCSR, lu, b, z=None, y=None):
        """L z = b (unit lower) then U y = z (non-unit upper) on the in-place LU values."""
        self._claim(A)
        z = torch.empty_like(b) if z is None else z
        y = torch.empty_like(b) if y is None else y
        fn = self.lib.sptrsv_lu_d if lu.dtype == torch.float64 else self.lib.sptrsv_lu_s
        _check(fn(self.h, A.m, A.stored, A.base, self._p(A.rowptr), self._p(A.colidx), self._p(lu), self._p(b),
                  self._p(z), self._p(y)), "sptrsv_lu")
        return z, y

    def bicgstab(self, A: DeviceCSR, b, x=None, lu=None, max_iters=1000, rtol=1e-8):
        """rsg_bicgstab_d: like pcg, for matrices that are not symmetric positive definite."""
        return self.pcg(A, b, x, lu, max_iters, rtol, _fn="rsg_bicgstab_d")

    def gmres(self, A: DeviceCSR, b, x=None, lu=None, max_iters=1000, rtol=1e-8):
        """rsg_gmres_d: restarted GMRES(30), right-preconditioned; an iteration is one Arnoldi step."""
        return self.pcg(A, b, x, lu, max_iters, rtol, _fn="rsg_gmres_d")

    def pcg(self, A: DeviceCSR, b, x=None, lu=None, max_iters=1000, rtol=1e-8, _fn="rsg_pcg_d"):
        """rsg_pcg_d: A (fp64 DeviceCSR) x = b, preconditioned by ILU(0) factors `lu` (fp64 or fp32 tensor, or None).
        Returns (x, iterations, relative residual)."""
        self._claim(A)
        assert A.precision == "d" and b.dtype == torch.float64
        x = torch.zeros_like(b) if x is None else x
        it, rel = C.c_int(0), C.c_double(0.0)
        prec = 0 if lu is None else (ord("d") if lu.dtype == torch.float64 else ord("s"))
        _check(getattr(self.lib, _fn)(self.h, A.m, A.stored, A.base, self._p(A.rowptr), self._p(A.colidx), self._p(A.values),
                                      self._p(lu), prec, self._p(b), self._p(x), int(max_iters), float(rtol), C.byref(it),
                                      C.byref(rel)), _fn)
        return x, it.value, rel.value

    def levels(self, which, n):
        """(level_ptr, level_rows) of the last analysed matrix; which 0 = L, 1 = U."""
        nl = C.c_int(0)
        _check(self.lib.rsg_get_levels(self.h, which, C.byref(nl), None, None), "rsg_get_levels")
        lp = np.zeros(nl.value + 1, dtype=np.int32)
        rows = np.zeros(max(n, 1), dtype=np.int32)
        _check(self.lib.rsg_get_levels(self.h, which, C.byref(nl), C.c_void_p(lp.ctypes.data), C.c_void_p(rows.ctypes.data)),
               "rsg_get_levels")
        return lp, rows[:n]

    def analyse_levels(self, A: DeviceCSR):
        """ILU0 analysis that tolerates a missing diagonal (levels are still defined)."""
        return self.ilu0_analysis(A)

    def row_partition(self, A: DeviceCSR, ngpu, balanced=True):
        self._claim(A)
        out = np.zeros(ngpu + 1, dtype=np.int32)
        _check(self.lib.rsg_get_row_partition(self.h, A.m, A.base, self._p(A.rowptr), ngpu, int(balanced),
                                              C.c_void_p(out.ctypes.data)), "rsg_get_row_partition")
        return out

    # ---- host-pointer entry points (numpy in, numpy out; H2D/D2H inside the call)
    def spmv_host(self, a: CSR, x, precision="d", alpha=1.0, beta=0.0, y=None):
        vt = np.float64 if precision == "d" else np.float32
        xt = np.float32 if precision == "s" else np.float64
        vals = np.ascontiguousarray(a.values, dtype=vt)
        x = np.ascontiguousarray(x, dtype=xt)
        y = np.zeros(a.m, dtype=xt) if y is None else np.ascontiguousarray(y, dtype=xt)
        fn = {"d": self.lib.spmv_csr_d_host, "s": self.lib.spmv_csr_s_host, "m": self.lib.spmv_csr_mixed_host}[precision]
        _check(fn(self.h, a.m, a.n, a.stored, a.base, a.rowptr.ctypes.data, a.colidx.ctypes.data, vals.ctypes.data,
                  alpha, x.ctypes.data, beta, y.ctypes.data), "spmv_csr_host")
        return y

    def ilu0_factor_host(self, a: CSR, dtype=np.float64):
        vals = np.array(a.values, dtype=dtype)
        zp = C.c_int(-1)
        fn = self.lib.ilu0_factor_d_host if dtype == np.float64 else self.lib.ilu0_factor_s_host
        st = fn(self.h, a.m, a.stored, a.base, a.rowptr.ctypes.data, a.colidx.ctypes.data, vals.ctypes.data, C.byref(zp))
        _check(st, "ilu0_factor_host", allow=(ZERO_PIVOT_STRUCTURAL, ZERO_PIVOT_NUMERIC))
        return st, zp.value, vals

    def sptrsv_lu_host(self, a: CSR, lu, b):
        lu = np.ascontiguousarray(lu)
        b = np.ascontiguousarray(b, dtype=lu.dtype)
        y = np.zeros(a.m, dtype=lu.dtype)
        fn = self.lib.sptrsv_lu_d_host if lu.dtype == np.float64 else self.lib.sptrsv_lu_s_host
        _check(fn(self.h, a.m, a.stored, a.base, a.rowptr.ctypes.data, a.colidx.ctypes.data, lu.ctypes.data,
                  b.ctypes.data, y.ctypes.data), "sptrsv_lu_host")
        return y

    def matrix_upload(self, a: CSR, precision="d"):
        mat = C.c_void_p()
        vals = np.ascontiguousarray(a.values, dtype=np.float64)
        _check(self.lib.rsg_matrix_upload(self.h, precision.encode(), a.m, a.n, a.stored, a.base, a.rowptr.ctypes.data,
                                          a.colidx.ctypes.data, vals.ctypes.data, C.byref(mat)), "rsg_matrix_upload")
        return mat

    def matrix_spmv_host(self, mat, x, y, alpha=1.0, beta=0.0):
        _check(self.lib.rsg_matrix_spmv_host(self.h, mat, alpha, x.ctypes.data, beta, y.ctypes.data), "rsg_matrix_spmv_host")
        return y

    def matrix_free(self, mat):
        _check(self.lib.rsg_matrix_free(self.h, mat), "rsg_matrix_free")
