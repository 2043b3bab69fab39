#!/usr/bin/env python
"""bench.py -- the driver's measurement contract for the ReSpaSol hot path on B200.

    python bench.py --gpus N --steps K --warmup W [--impl reference]

Step   = one CSR SpMV (the hot path, BASELINE.json configs[1]: 3-D 27-point stencil 100^3, 1M rows,
         26.46M entries, fp64) over one of R resident replicas of the matrix (rotated so that the
         ~338 MB a step streams never comes out of the 126 MB L2: "inputs larger than L2").
Metric = achieved HBM GB/s on ALGORITHMIC bytes (12*nnz + 4(m+1) + 8m + 8n for fp64).
N > 1  = weak scaling: every rank owns a 100x100x100-plane slab of a 100x100x(100N) grid; a step is
         the x-halo all-gather (NCCL over NVLink) followed by the local SpMV.
The line also carries fp32 / mixed SpMV, the skewed (merge-path) config, ILU0 analysis/factor/solve
milliseconds (config 3), the end-to-end figure through the host-pointer C ABI, the roofline object,
the host-CPU baseline and the clocks seen during the timed region.
--impl reference times the reference's CPU implementation of the path (its MKL sparse mv is absent
from this image: the oracle's OpenMP row-parallel restatement stands in) on all host threads.
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import numpy as np  # noqa: E402

GRID = 100          # configs[1]: 27-point stencil on 100^3
REPLICAS = 2
METRIC = "SpMV achieved HBM GB/s (fp64 CSR, algorithmic bytes / device time)"


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def gpu_local_cpus(index):
    """CPUs on the NUMA node the GPU hangs off (NVML's ideal affinity), or None.  A page-locked buffer that lives on
    the other socket costs the H2D copy up to 2x (measured on this pool: 8 MB in 157 us vs 287 us), so the end-to-end
    leg allocates its host buffers and issues its copies from these CPUs -- what `numactl` would do for the harness."""
    try:
        import pynvml
        pynvml.nvmlInit()
        dev = pynvml.nvmlDeviceGetHandleByIndex(index)
        words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(dev, words)
        cpus = {64 * w + b for w, m in enumerate(mask) for b in range(64) if (m >> b) & 1}
        cpus &= os.sched_getaffinity(0)
        return cpus or None
    except Exception:  # noqa: BLE001
        return None


class ClockSampler(threading.Thread):
    """nvidia-smi's clocks / throttle reasons through NVML, polled while the timed region runs."""

    def __init__(self, index, period=0.02):
        super().__init__(daemon=True)
        self.index, self.period = index, period
        self.samples, self.stop_flag, self.ok = [], False, False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.dev = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_sm = pynvml.nvmlDeviceGetMaxClockInfo(self.dev, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception as e:  # noqa: BLE001
            self.err = repr(e)

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        while not self.stop_flag:
            try:
                sm = nv.nvmlDeviceGetClockInfo(self.dev, nv.NVML_CLOCK_SM)
                rs = nv.nvmlDeviceGetCurrentClocksEventReasons(self.dev) if hasattr(nv, "nvmlDeviceGetCurrentClocksEventReasons") \
                    else nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.dev)
                pw = nv.nvmlDeviceGetPowerUsage(self.dev) / 1000.0
                self.samples.append((time.time(), sm, rs, pw))
            except Exception:  # noqa: BLE001
                pass
            time.sleep(self.period)

    def summary(self, t0, t1):
        if not self.ok:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvml unavailable: " + getattr(self, "err", "?")]}
        sel = [s for s in self.samples if t0 <= s[0] <= t1] or self.samples[-3:]
        names = {0x2: "applications_clocks_setting", 0x4: "sw_power_cap", 0x8: "hw_slowdown", 0x10: "sync_boost",
                 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown", 0x80: "hw_power_brake_slowdown",
                 0x100: "display_clock_setting"}
        bits = 0
        for s in sel:
            bits |= s[2]
        reasons = [n for b, n in names.items() if bits & b]
        return {"sm_mhz": float(np.median([s[1] for s in sel])) if sel else None, "sm_max_mhz": float(self.max_sm),
                "reasons": reasons, "samples": len(sel), "power_w_max": max([s[3] for s in sel]) if sel else None}


def spmv_bytes(stored, m, n, precision):
    v, x = {"d": (8, 8), "s": (4, 4), "m": (4, 8)}[precision]
    return stored * (v + 4) + (m + 1) * 4 + m * x + n * x


# ------------------------------------------------------------------------------------ reference arm
def run_reference(args):
    """The reference's CPU SpMV path (test_spmv.c:168-180, mkl_sparse_d_mv under OpenMP) on this box's
    host cores.  MKL is not in the image and cannot be installed: the oracle's OpenMP restatement of
    that call is timed instead (kind "port").  Rank 0 only."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    import orc
    from respasol_b200 import host
    threads = orc.max_threads()
    a = host.gen_stencil27(GRID, seed=2)
    x = orc.dlarnv(a.n) - 0.5
    bytes_per = spmv_bytes(a.stored, a.m, a.n, "d")
    for _ in range(max(args.warmup, 1)):
        orc.spmv(a, x, "d", threads=threads)
    steps = max(1, min(args.steps, 200))  # bounded sample: each product is tens of ms on the host
    t0 = time.perf_counter()
    for _ in range(steps):
        y = orc.spmv(a, x, "d", threads=threads)
    dt = time.perf_counter() - t0
    val = bytes_per * steps / dt * 1e-9
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": "GB/s", "n_gpus": args.gpus, "steps": steps,
            "warmup": max(args.warmup, 1), "ms_per_step": dt / steps * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "CSR SpMV fp64, 3-D 27-point stencil 100^3 (1M rows, 26463592 entries), host CPU"},
            "cpu_baseline": {"value": val, "unit": "GB/s", "cores": threads, "kind": "port",
                             "sample": "%d products of the full config-2 matrix, OpenMP static rows, all %d host threads; "
                                       "stands in for mkl_sparse_d_mv (MKL absent)" % (steps, threads)},
            "e2e": {"value": val, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gflops": 2.0 * a.stored * steps / dt * 1e-9, "checksum": float(np.abs(y).sum())}
    print(json.dumps(line))


# ------------------------------------------------------------------------------------ our arm
def timed_loop(torch, fn, steps, barrier):
    """K calls bracketed by barrier + synchronize, timed with CUDA events on the launching stream."""
    start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    torch.cuda.synchronize()
    w0 = time.time()
    start.record()
    for i in range(steps):
        fn(i)
    stop.record()
    torch.cuda.synchronize()
    w1 = time.time()
    barrier()
    return start.elapsed_time(stop), w0, w1  # ms


def run_ours(args):
    import torch
    import torch.distributed as dist
    import orc
    from respasol_b200 import host
    from respasol_b200.dist import RowBlockSpmv
    from respasol_b200.gpu import ALGO, DeviceCSR, Handle

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback)"
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        # the box sets NCCL_DEBUG=VERSION, which prints a banner on STDOUT ahead of the JSON line: keep stdout to the one line
        if os.environ.get("NCCL_DEBUG", "").upper() == "VERSION":
            del os.environ["NCCL_DEBUG"]  # WARN and above print the banner too
        dist.init_process_group("nccl", device_id=dev)
    assert world == args.gpus or world == 1 and args.gpus == 1, "launch with torchrun --nproc-per-node N for --gpus N"

    def barrier():
        if world > 1:
            dist.barrier()

    def allmax(v):
        if world == 1:
            return v
        t = torch.tensor([v], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def allsum(v):
        if world == 1:
            return v
        t = torch.tensor([v], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    h = Handle(local_rank)
    K, W = args.steps, max(args.warmup, 3)
    peak, peak_src = peaks()
    sampler = ClockSampler(local_rank)
    sampler.start()
    extra = {}

    # ---- the timed workload: this rank's slab of the 100 x 100 x (100*world) stencil
    nz = GRID * world
    a = host.gen_stencil27(GRID, GRID, nz, rank * GRID, (rank + 1) * GRID, seed=2)
    bytes_per = spmv_bytes(a.stored, a.m, a.m, "d")  # per rank: its rows, its slice of x
    xs = orc.dlarnv(GRID * GRID * nz)[rank * a.m:(rank + 1) * a.m] - 0.5
    launches0 = h.info("KERNEL_LAUNCHES")
    if world == 1:
        mats = [DeviceCSR(a, "d", dev) for _ in range(REPLICAS)]
        xd = [torch.from_numpy(xs).to(dev) for _ in range(REPLICAS)]
        yd = [torch.empty(a.m, dtype=torch.float64, device=dev) for _ in range(REPLICAS)]

        def step(i):
            r = i % REPLICAS
            h.spmv(mats[r], xd[r], yd[r])
        for r in range(REPLICAS):
            h.spmv_analysis(mats[r])
        launches_per_step = 1
    else:
        row_starts = np.arange(world + 1, dtype=np.int64) * a.m
        ops = []
        hp = Handle(local_rank)  # the push kernels run on a stream of their own, beside the product
        side = torch.cuda.Stream(device=dev)
        with torch.cuda.stream(side):
            hp.use_current_stream()
        if args.exchange == "p2p-barrier":
            hp.use_current_stream()
        push = hp.halo_push
        P2P_FNS = {"p2p": (push, h.spmv_halo_wait, h.spmv_halo_fuse), "p2p-side": (push, h.spmv_halo_wait),
                   "p2p-barrier": (push, None), "nccl": (push, None)}[args.exchange]
        for r in range(REPLICAS):
            op = RowBlockSpmv(a, row_starts, "p2p" if args.exchange.startswith("p2p") else "halo", lambda c: DeviceCSR(c, "d", dev),
                              lambda A, x, y: h.spmv(A, x, y), dev, torch.float64,
                              view=(lambda A, r0, r1: A.view_rows(r0, r1)) if args.overlap else None,
                              push=P2P_FNS, side=side)
            op.set_x(torch.from_numpy(xs).to(dev))
            if op.p2p is not None and op.p2p[1] is not None:
                h.set_option("SPMV_HALO_START", op.plan.halo_base)  # marks the tiles that must wait for the peers
            for A in ([piece for piece, _, _ in op.split] if op.split else [op.A]):
                h.spmv_analysis(A)
            ops.append(op)
        yd = [op.y for op in ops]
        if args.graph:
            for op in ops:
                for _ in range(3):  # eager steps first: plans, NCCL channels, allocator all settled
                    op.step()
                op.capture(h.use_current_stream)
            extra["cuda_graph"] = "one graph per step (pack + NCCL all-gather + SpMV launches), replayed"

        def step(i):
            ops[i % REPLICAS].step()
        launches_per_step = len(ops[0].split) if ops[0].split else 1
        launches_per_step += 1 if (ops[0].p2p is not None and len(ops[0].p2p) < 3) else 0
        extra["exchange"] = {"mode": (("halo push over NVLink peer memory (rsg_halo_push) + " +
                                       ("FUSED into the SpMV launch (first CTAs push + signal, boundary tiles await the neighbours' flags)" if len(ops[0].p2p) > 2
                                        else "arrival flags awaited inside the SpMV's boundary tiles" if ops[0].p2p[1] else "symmetric-memory barrier"))
                                      if ops[0].p2p is not None else "halo all-gather (NCCL)") + (" on a side stream, overlapped with the %d interior rows; "
                                                                   "boundary rows follow" % ops[0].interior_rows if ops[0].split else ""),
                             "bytes_received_per_rank_per_step": ops[0].exchange_bytes}

    for i in range(W):
        step(i)
    launches_before = h.info("KERNEL_LAUNCHES")
    push_launches_before = hp.info("KERNEL_LAUNCHES") if world > 1 else 0
    ms, w0, w1 = timed_loop(torch, step, K, barrier)
    gpu_launches = h.info("KERNEL_LAUNCHES") - launches_before
    if world > 1:
        gpu_launches += hp.info("KERNEL_LAUNCHES") - push_launches_before
    if "cuda_graph" in extra:  # replayed launches do not pass through the library's counter
        gpu_launches = K * launches_per_step
    if world > 1:  # how long the host needs to issue a step (a step cannot be faster than this)
        torch.cuda.synchronize()
        c0 = time.perf_counter()
        for i in range(200):
            step(i)
        extra["host_issue_us_per_step"] = (time.perf_counter() - c0) / 200 * 1e6
        torch.cuda.synchronize()
    ms = allmax(ms)
    total_bytes = allsum(float(bytes_per))
    value = total_bytes * K / (ms * 1e-3) * 1e-9
    # keep the load up a little longer so the 20 ms clock poll sees the kernel even for a short K
    t_end = time.time() + max(0.0, 0.4 - (w1 - w0))
    while time.time() < t_end:
        for i in range(200):
            step(i)
        torch.cuda.synchronize()
    w1b = time.time()
    clocks = sampler.summary(w0, max(w1, w1b))

    h.set_option("SPMV_HALO_START", 0)
    # parity spot check of what was just timed (first rows of this rank against the oracle)
    if world == 1:
        ref = orc.spmv(a.rows(0, 2000), xs, "d")
        got = yd[0][:2000].cpu().numpy()
        assert np.abs(got - ref).max() <= 1e-12 * np.abs(ref).max(), "bench result differs from the oracle"
    else:
        # both ends of this rank's block read the neighbours' entries: they check the exchange as well
        xg = orc.dlarnv(GRID * GRID * nz) - 0.5
        for r0, r1 in ((0, 3000), (a.m - 3000, a.m)):
            ref = orc.spmv(a.rows(r0, r1), xg, "d")
            for yv in yd:
                got = yv[r0:r1].cpu().numpy()
                assert np.abs(got - ref).max() <= 1e-12 * np.abs(ref).max(), "rank %d: distributed result differs from the oracle" % rank
        extra["parity"] = "rows at both ends of every rank's block (they read the neighbours' halo) match the oracle to 1e-12"

    # ---- roofline of the dominant kernel (single GPU: the step IS one launch of it)
    per_launch_ms = ms / K
    achieved = bytes_per / (per_launch_ms * 1e-3) * 1e-9
    traffic = None
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tp):
        with open(tp) as f:
            traffic = json.load(f).get("spmv_tile_fp64_c2_bytes_per_launch")
    roofline = {"bound": "hbm", "kernel": "spmv_tile_kernel<double,double,128,2304,512>", "achieved": achieved, "peak": peak,
                "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": bytes_per, "frac_of_8TBs_nominal": achieved / 8000.0,
                "us_per_launch": per_launch_ms * 1e3}

    # ---- BASELINE configs[4]: the 64M-entry matrix (27-point, 134^3), STRONG scaling: plane-aligned row blocks
    # over all ranks, same exchange as above; two resident copies rotated (per-rank share > L2 only up to N=4,
    # so the copies keep a step from being served out of L2)
    if not args.no_c5:
        G5 = 134
        cuts = [(r * G5 + world // 2) // world for r in range(world + 1)]
        a5 = host.gen_stencil27(G5, G5, G5, cuts[rank], cuts[rank + 1], seed=2)
        starts5 = np.array(cuts, dtype=np.int64) * G5 * G5
        x5 = (orc.dlarnv(G5 ** 3) - 0.5)[starts5[rank]:starts5[rank + 1]]
        ops5 = []
        for r in range(REPLICAS):
            if world > 1:
                op = RowBlockSpmv(a5, starts5, "p2p" if args.exchange.startswith("p2p") else "halo", lambda c: DeviceCSR(c, "d", dev),
                                  lambda A, x, y: h.spmv(A, x, y), dev, torch.float64,
                                  push=P2P_FNS, side=side)
                if op.p2p is not None and op.p2p[1] is not None:
                    h.set_option("SPMV_HALO_START", op.plan.halo_base)
            else:
                op = RowBlockSpmv(a5, starts5, "halo", lambda c: DeviceCSR(c, "d", dev), lambda A, x, y: h.spmv(A, x, y), dev,
                                  torch.float64)
            op.set_x(torch.from_numpy(x5.copy()).to(dev))
            h.spmv_analysis(op.A)
            ops5.append(op)
        for i in range(6):
            ops5[i % REPLICAS].step()
        K5 = max(50, min(K, 400))
        ms5, _, _ = timed_loop(torch, lambda i: ops5[i % REPLICAS].step(), K5, barrier)
        ms5 = allmax(ms5)
        stored5 = allsum(float(a5.stored))
        b5 = stored5 * 12 + (G5 ** 3 + 1) * 4 + 2 * 8 * G5 ** 3
        ref5 = orc.spmv(a5.rows(a5.m - 2000, a5.m), orc.dlarnv(G5 ** 3) - 0.5, "d")
        got5 = ops5[(K5 - 1) % REPLICAS].y[a5.m - 2000:].cpu().numpy()
        assert np.abs(got5 - ref5).max() <= 1e-12 * np.abs(ref5).max(), "rank %d: config-5 result differs from the oracle" % rank
        extra["c5_strong_fp64"] = {"workload": "27-point 134^3, %d entries in total, row blocks of whole planes over %d GPU(s)" % (int(stored5), world),
                                   "us_per_product": ms5 / K5 * 1e3, "GBps_aggregate": b5 / (ms5 / K5 * 1e-3) * 1e-9,
                                   "GFLOPs": 2.0 * stored5 / (ms5 / K5 * 1e-3) * 1e-9, "scaling": "strong"}
        h.set_option("SPMV_HALO_START", 0)
        del ops5, a5

    e2e = None
    cpu_baseline = None
    full = a if world == 1 else host.gen_stencil27(GRID, seed=2)
    xfull = orc.dlarnv(full.n) - 0.5
    if world == 1:
        # ---- other precisions / configs (single-GPU kernels; reported, not the headline)
        def rate(A, x, y, reps):
            for _ in range(3):
                h.spmv(A, x, y)
            t, _, _ = timed_loop(torch, lambda i: h.spmv(A, x, y), reps, lambda: None)
            return t / reps

        reps = max(20, min(K, 500))
        for p in ("s", "m"):
            As = [DeviceCSR(full, p, dev) for _ in range(REPLICAS)]
            xt = torch.float32 if p == "s" else torch.float64
            xv = torch.from_numpy(xfull).to(dev).to(xt)
            yv = torch.empty(full.m, dtype=xt, device=dev)
            for _ in range(3):
                for A in As:
                    h.spmv(A, xv, yv)
            t, _, _ = timed_loop(torch, lambda i: h.spmv(As[i % REPLICAS], xv, yv), reps, lambda: None)
            b = spmv_bytes(full.stored, full.m, full.n, p)
            extra["spmv_" + {"s": "fp32", "m": "mixed"}[p]] = {"GBps": b / (t / reps * 1e-3) * 1e-9, "us": t / reps * 1e3,
                                                                "GFLOPs": 2.0 * full.stored / (t / reps * 1e-3) * 1e-9}
            del As
        extra["spmv_fp64"] = {"GBps": achieved, "us": per_launch_ms * 1e3, "GFLOPs": 2.0 * a.stored / (per_launch_ms * 1e-3) * 1e-9}
        # config 4: skewed rows -- tile kernel with its long rows split into segment CTAs (fits L2: warm figure, labelled)
        c4 = host.gen_powerlaw(320000, 2.2, 64, 4, 100000, 4)
        A4 = DeviceCSR(c4, "d", dev)
        x4 = torch.from_numpy(orc.dlarnv(c4.n)).to(dev)
        y4 = torch.empty(c4.m, dtype=torch.float64, device=dev)
        t4 = rate(A4, x4, y4, reps)
        extra["spmv_c4_powerlaw_fp64"] = {"us": t4 * 1e3, "GBps_L2_warm": spmv_bytes(c4.stored, c4.m, c4.n, "d") / (t4 * 1e-3) * 1e-9,
                                          "algo": h.info("SPMV_ALGO_USED"), "max_row": h.info("SPMV_MAX_ROW"), "stored": c4.stored}
        h.set_option("SPMV_ALGO", ALGO["rowblock"])
        t4r = rate(A4, x4, y4, max(5, reps // 10))
        h.set_option("SPMV_ALGO", 0)
        extra["spmv_c4_powerlaw_fp64"]["us_rowblock_cta_per_row"] = t4r * 1e3
        h.set_option("SPMV_ALGO", ALGO["merge"])
        t4m = rate(A4, x4, y4, max(5, reps // 10))
        h.set_option("SPMV_ALGO", 0)
        extra["spmv_c4_powerlaw_fp64"]["us_merge_path"] = t4m * 1e3
        # config 1 (fits L2: warm figure, labelled)
        c1 = host.gen_laplace2d(1000)
        A1 = DeviceCSR(c1, "d", dev)
        x1 = torch.ones(c1.n, dtype=torch.float64, device=dev)
        y1 = torch.empty(c1.m, dtype=torch.float64, device=dev)
        t1 = rate(A1, x1, y1, reps)
        extra["spmv_c1_lap2d_fp64"] = {"us": t1 * 1e3, "GBps_L2_warm": spmv_bytes(c1.stored, c1.m, c1.n, "d") / (t1 * 1e-3) * 1e-9}
        # config 3: ILU0 analysis / factor / L+U solve, median of 10
        c3 = host.gen_femlike(47, 0.58, 3)
        for dt_, p in ((torch.float64, "d"), (torch.float32, "s")):
            A3 = DeviceCSR(c3, p, dev)
            pristine = A3.values.clone()
            b3 = torch.ones(c3.m, dtype=dt_, device=dev)
            z3, y3 = torch.empty_like(b3), torch.empty_like(b3)
            ta, tf, ts = [], [], []
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
            for it in range(11):
                A3.values.copy_(pristine)
                torch.cuda.synchronize()
                ev[0].record(); h.ilu0_analysis(A3); h.sptrsv_analysis(A3); ev[1].record(); torch.cuda.synchronize()
                ta.append(ev[0].elapsed_time(ev[1]))
                ev[0].record(); h.ilu0_factor(A3, report_pivot=False); ev[1].record(); torch.cuda.synchronize()
                tf.append(ev[0].elapsed_time(ev[1]))
                ev[0].record(); h.sptrsv_lu(A3, A3.values, b3, z3, y3); ev[1].record(); torch.cuda.synchronize()
                ts.append(ev[0].elapsed_time(ev[1]))
            extra["ilu0_c3_" + {"d": "fp64", "s": "fp32"}[p]] = {
                "analysis_ms": float(np.median(ta[1:])), "factor_ms": float(np.median(tf[1:])),
                "solve_LU_ms": float(np.median(ts[1:])), "levels_L": h.info("LEVELS_L"), "levels_U": h.info("LEVELS_U"),
                "n": c3.m, "stored": c3.stored}
        # the loop around the path: ILU(0)-preconditioned BiCGStab and flexible GMRES(30) on config 3, fp64 and fp32 factors
        A3 = DeviceCSR(c3, "d", dev)
        xs3 = torch.from_numpy(orc.dlarnv(c3.n) - 0.5).to(dev)
        rhs3 = h.spmv(A3, xs3).clone()
        for p in ("d", "s"):
            F3 = DeviceCSR(c3, p, dev)
            h.ilu0_factor(F3, report_pivot=False)
            for solver in ("bicgstab", "gmres"):
                run = getattr(h, solver)
                run(A3, rhs3, lu=F3.values, max_iters=200, rtol=1e-10)
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                xk, itk, relk = run(A3, rhs3, lu=F3.values, max_iters=200, rtol=1e-10)
                torch.cuda.synchronize()
                extra[solver + "_ilu0_c3_" + {"d": "fp64_factors", "s": "fp32_factors"}[p]] = {
                    "iterations": itk, "relres": relk, "ms": (time.perf_counter() - t0) * 1e3,
                    "max_err": float((xk - xs3).abs().max())}
    # ---- end to end through the host-pointer C ABI (matrix resident like the harness, spmv.cu:111-118;
    # every step: x from pinned host memory -> device, SpMV, y device -> host).  Every rank runs its own
    # config-2 product; the aggregate is all ranks' bytes over the slowest rank's time.
    full64 = full
    all_cpus = os.sched_getaffinity(0)
    near = gpu_local_cpus(dev.index if dev.index is not None else 0)
    if near:
        os.sched_setaffinity(0, near)
    mat = h.matrix_upload(full64, "d")
    # inputs and result live in page-locked host memory (torch pinned tensors viewed as numpy)
    xh_t = torch.from_numpy(np.ascontiguousarray(xfull)).pin_memory()
    yh_t = torch.zeros(full64.m, dtype=torch.float64).pin_memory()
    xh, yh = xh_t.numpy(), yh_t.numpy()
    def pcie_link():
        try:
            import pynvml
            pynvml.nvmlInit()
            d = pynvml.nvmlDeviceGetHandleByIndex(dev.index if dev.index is not None else 0)
            return "gen%d x%d" % (pynvml.nvmlDeviceGetCurrPcieLinkGeneration(d), pynvml.nvmlDeviceGetCurrPcieLinkWidth(d))
        except Exception:  # noqa: BLE001
            return "unknown"

    def copy_alone_us(dst, src, reps=20):
        for _ in range(3):
            dst.copy_(src, non_blocking=True)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(reps):
            dst.copy_(src, non_blocking=True)
            torch.cuda.synchronize()
        return (time.perf_counter() - t0) / reps * 1e6

    # the e2e products keep the SMs almost idle; make sure the device is out of its low-power state first
    # (0.3 s of device-resident products), then look at what the PCIe link does on its own on this box
    spin_x = torch.from_numpy(xfull).to(dev)
    spin_y = torch.empty(full64.m, dtype=torch.float64, device=dev)
    spin_A = DeviceCSR(full64, "d", dev)
    t0 = time.perf_counter()
    while time.perf_counter() - t0 < 0.3:
        for _ in range(50):
            h.spmv(spin_A, spin_x, spin_y)
        torch.cuda.synchronize()
    h2d_alone = copy_alone_us(spin_x, xh_t)
    d2h_alone = copy_alone_us(yh_t, spin_y)
    del spin_A
    for _ in range(3):
        h.matrix_spmv_host(mat, xh, yh)
    e_steps = max(50, min(K, 200))
    # PCIe time on a shared host is noisy from run to run (8 MB H2D measured between 157 and 287 us on this pool):
    # five rounds of e_steps products, the median round is reported, the spread is kept
    rounds = []
    links = []
    for _ in range(5):
        barrier()
        t0 = time.perf_counter()
        for _ in range(e_steps):
            h.matrix_spmv_host(mat, xh, yh)
        rounds.append(allmax((time.perf_counter() - t0) / e_steps))
        links.append(pcie_link())
    edt = sorted(rounds)[len(rounds) // 2]
    h.matrix_free(mat)
    os.sched_setaffinity(0, all_cpus)
    refy = orc.spmv(full64.rows(0, 1000), xfull, "d")
    assert np.abs(yh[:1000] - refy).max() <= 1e-12 * np.abs(refy).max()
    e2e = {"value": world * spmv_bytes(full64.stored, full64.m, full64.n, "d") / edt * 1e-9, "unit": "GB/s",
           "h2d_bytes_per_step": 8 * full64.n * world, "d2h_bytes_per_step": 8 * full64.m * world, "ms_per_step": edt * 1e3,
           "host_cpus": ("%d GPU-local of %d" % (len(near), len(all_cpus))) if near else "unbound",
           "ms_per_step_rounds": [round(r * 1e3, 4) for r in rounds],
           "pcie": {"link_during_rounds": sorted(set(links)), "h2d_alone_us": round(h2d_alone, 1), "d2h_alone_us": round(d2h_alone, 1),
                    "note": "plain cudaMemcpyAsync of the same 8 MB buffers on this box, measured just before the rounds"},
           "note": "rsg_matrix_spmv_host on every rank: matrix resident in HBM (uploaded once, as GPU/spmv.cu:111-118 does), "
                   "x H2D (from pinned host memory) + SpMV + y D2H (to pinned host memory) inside every timed step"}
    if world == 1:
        # ---- CPU baseline on this box's host cores (bounded sample)
        if not args.no_cpu:
            threads = orc.max_threads()
            orc.spmv(full64, xfull, "d", threads=threads)
            n_cpu, t0 = 0, time.perf_counter()
            while time.perf_counter() - t0 < 8.0 and n_cpu < 200:
                orc.spmv(full64, xfull, "d", threads=threads)
                n_cpu += 1
            cdt = (time.perf_counter() - t0) / n_cpu
            cpu_baseline = {"value": spmv_bytes(full64.stored, full64.m, full64.n, "d") / cdt * 1e-9, "unit": "GB/s",
                            "cores": threads, "kind": "port", "ms_per_product": cdt * 1e3,
                            "sample": "%d fp64 products of the full config-2 matrix (~8 s), OpenMP static rows on all host "
                                      "threads; stands in for mkl_sparse_d_mv (test_spmv.c:168-170), MKL absent" % n_cpu}
            t0 = time.perf_counter()
            orc.ilu0(c3)
            extra["cpu_ilu0_c3_fp64_ms"] = (time.perf_counter() - t0) * 1e3
    sampler.stop_flag = True
    barrier()
    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": "GB/s", "n_gpus": world, "steps": K, "warmup": W,
                "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
                "data": "synthetic",
                "config": {"workload": "CSR SpMV fp64, 3-D 27-point stencil 100x100x%d, %d rows / %d entries per GPU "
                                       "(BASELINE configs[1] per GPU)" % (nz, a.m, a.stored),
                           "l2_policy": "inputs larger than L2: %d resident replicas of the matrix and vectors rotated "
                                        "(%.0f MB per step, %.0f MB working set vs 126 MB L2)"
                                        % (REPLICAS, bytes_per / 1e6, REPLICAS * bytes_per / 1e6),
                           "parallelism": "row-block x%d, x-halo all-gather" % world if world > 1 else "single GPU"},
                "gflops": 2.0 * a.stored * world * K / (ms * 1e-3) * 1e-9,
                "frac_of_8TBs_nominal": value / (8000.0 * world),
                "clocks": clocks, "e2e": e2e, "gpu_launches": int(gpu_launches), "launches_per_step": launches_per_step,
                "roofline": roofline, "cpu_baseline": cpu_baseline, "extra": extra}
        print(json.dumps(line))
    if world > 1:
        # symmetric-memory handles / NCCL communicators are torn down by process exit: an orderly
        # destroy_process_group after P2P mappings (or captured graphs) can block on a peer that is already gone
        dist.barrier()
        torch.cuda.synchronize()
        sys.stdout.flush()
        os._exit(0)
    h.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the host-CPU baseline leg")
    ap.add_argument("--no-c5", action="store_true", help="skip the 64M-entry strong-scaling leg (BASELINE configs[4])")
    ap.add_argument("--graph", action="store_true", help="N>1: replay each step from a CUDA graph (NCCL exchange only)")
    ap.add_argument("--overlap", action="store_true", help="N>1: exchange on a side stream while the interior rows are multiplied")
    ap.add_argument("--exchange", default="p2p", choices=["p2p", "p2p-side", "p2p-barrier", "nccl"], help="N>1: x-halo exchange: peer-memory push or NCCL all-gather")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
