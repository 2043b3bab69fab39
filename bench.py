#!/usr/bin/env python
"""bench.py -- the driver's measurement contract for the ReSpaSol hot path on B200.

    python bench.py --gpus N --steps K --warmup W [--impl reference]

Step   = one CSR SpMV (the hot path) over one of R resident replicas of the matrix, rotated so that what a step
         streams never comes out of L2 ("inputs larger than L2").
N = 1  : BASELINE.json configs[1] -- 3-D 27-point stencil 100^3, 1M rows, 26.46M entries, fp64.
N > 1  : BASELINE.json configs[4] -- the 64M-entry matrix (27-point 134^3), STRONG scaling: plane-aligned row blocks
         over the N ranks, x halo pushed over NVLink peer memory from inside the product's launch.  The weak-scaled
         configs[1]-per-GPU curve of round 1 is kept in `extra`.
Metric = achieved HBM GB/s on ALGORITHMIC bytes (12*nnz + 4(m+1) + 8m + 8n for fp64).  The x-window
         kernel reads LESS than that from HBM (16-bit window-local column indices built at plan time: 10 instead of
         12 bytes per entry) -- `roofline.traffic` and `roofline.dram_GBps` state what really moved.
The line also carries fp32 / mixed SpMV, cold-L2 configs[0] / configs[3], ILU0 analysis / factor / solve milliseconds
(configs[2]), cuSPARSE on the same box under one protocol, the end-to-end figure through the host-pointer C ABI, the
roofline object, the host-CPU baseline (all cores, and the reference's own protocol: 4 pinned threads, cold first
call, fp32, CPU ILU0 + L/U solves) and the clocks seen during the timed region.
--impl reference times the reference's CPU implementation of the path (its MKL sparse mv is absent from this image:
the oracle's OpenMP row-parallel restatement stands in) on all host cores of the box, rank 0 only.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import numpy as np  # noqa: E402

GRID = 100          # configs[1]: 27-point stencil on 100^3
GRID5 = 134         # configs[4]: 27-point stencil on 134^3 = 64 000 000 entries
REPLICAS = 2
METRIC = "SpMV achieved HBM GB/s (fp64 CSR, algorithmic bytes / device time)"
CHAIN_ALPHA = 1.0 / 52   # N > 1: x <- A x / 52 keeps the iterates O(1) (the stencil's rows sum to at most 52 in magnitude)
CHAIN_PARITY_STEPS = 12
LAUNCH_NOTE = ("back-to-back products of a static matrix, as in the reference's timed loop (GPU/spmv.cu:174-207); programmatic dependent "
               "launch with the values promised static (RSG_OPT_SPMV_PDL = 2) unless --pdl says otherwise")


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def job_config(n_gpus):
    """The `config` object: identical in both arms (ours / --impl reference) for a given N."""
    if n_gpus == 1:
        return {"workload": "CSR SpMV fp64, 3-D 27-point stencil 100^3 (1000000 rows, 26463592 entries) -- BASELINE configs[1]",
                "l2_policy": "inputs larger than L2: %d resident replicas of the matrix and vectors are rotated" % REPLICAS,
                "parallelism": "single GPU", "launch": LAUNCH_NOTE}
    return {"workload": "CSR SpMV fp64, 3-D 27-point stencil 134^3 (2406104 rows, 64000000 entries) -- BASELINE configs[4], strong scaling",
            "l2_policy": "inputs larger than L2 up to N = 4; %d resident replicas of every rank's block are rotated at every N" % REPLICAS,
            "parallelism": "row blocks of whole planes over %d GPUs; x halo pushed over NVLink peer memory by the product's own launch" % n_gpus,
            "launch": LAUNCH_NOTE}


def all_host_cpus():
    try:
        return sorted(os.sched_getaffinity(0))
    except AttributeError:
        return list(range(os.cpu_count() or 1))


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


def as_f32(a):
    """the same CSR with fp32 values (so that the timed CPU product does not cast 26M values per call)"""
    from respasol_b200 import host
    return host.CSR(a.m, a.n, a.nnz, a.rowptr, a.colidx, a.values.astype(np.float32), a.base)


# ------------------------------------------------------------------------------------ host-CPU legs
def cpu_spmv_rate(orc, a, x, precision, threads, seconds, max_products):
    """(products, seconds per product) of the oracle's OpenMP row-parallel SpMV: one warm-up, then a bounded sample"""
    orc.spmv(a, x, precision, threads=threads)
    n, t0 = 0, time.perf_counter()
    while (time.perf_counter() - t0 < seconds and n < max_products) or n == 0:
        y = orc.spmv(a, x, precision, threads=threads)
        n += 1
    return n, (time.perf_counter() - t0) / n, y


def cpu_leg_child(which):
    """Runs in a FRESH process (affinity and OMP_NUM_THREADS set by the parent before the OpenMP runtime exists).
    pinned4: the reference's own protocol -- OMP_NUM_THREADS=4 taskset -c 0,1,2,3 (run_spmv.sh:45), the single cold first
    call (test_spmv.c:168-170) next to the median of 50 warm calls, fp64 and fp32 (test_spmv.c:177-183);
    ilu: sequential IKJ ILU0 + L and U solves on configs[2] (stand-in for SuperLU dgsisx, test_superILU.c:118-124) and,
    labelled non-equivalent, scipy's spilu (SuperLU ILUTP, the algorithm dgsisx runs)."""
    import orc
    from respasol_b200 import host
    out = {}
    if which == "pinned4":
        a = host.gen_stencil27(GRID, seed=2)
        x = orc.dlarnv(a.n)  # U(0,1) like LAPACKE_dlarnv(1, ...) at test_spmv.c:75-76
        threads = int(os.environ.get("OMP_NUM_THREADS", "4"))
        for p, mat, xv in (("d", a, x), ("s", as_f32(a), x.astype(np.float32))):
            t0 = time.perf_counter()
            orc.spmv(mat, xv, p, threads=threads)  # the un-warmed call the reference times
            cold = time.perf_counter() - t0
            ts = []
            for _ in range(50):
                t0 = time.perf_counter()
                orc.spmv(mat, xv, p, threads=threads)
                ts.append(time.perf_counter() - t0)
            med = float(np.median(ts))
            b = spmv_bytes(a.stored, a.m, a.n, p)
            out["fp64" if p == "d" else "fp32"] = {"cold_first_call_ms": cold * 1e3, "warm_median_ms": med * 1e3,
                                                    "GBps_warm": b / med * 1e-9, "GBps_cold": b / cold * 1e-9}
        out["threads"] = threads
        out["cpus"] = sorted(os.sched_getaffinity(0))
        out["protocol"] = "OMP_NUM_THREADS=4 taskset -c 0,1,2,3 (run_spmv.sh:45); first call un-warmed (test_spmv.c:168-170), median of 50 warm calls"
    elif which == "ilu":
        c3 = host.gen_femlike(47, 0.58, 3)
        tf, ts = [], []
        for _ in range(3):
            t0 = time.perf_counter()
            st, piv, lu = orc.ilu0(c3)
            tf.append(time.perf_counter() - t0)
            b = np.ones(c3.m)
            t0 = time.perf_counter()
            orc.sptrsv_lu(c3, lu, b)
            ts.append(time.perf_counter() - t0)
        out["cpu_ilu0_c3_fp64"] = {"factor_ms": float(np.median(tf)) * 1e3, "solve_LU_ms": float(np.median(ts)) * 1e3, "threads": 1,
                                   "label": "CPU ILU0 restatement (sequential IKJ + sequential L/U solves); the reference used SuperLU "
                                            "ILUTP dgsisx (test_superILU.c:118-124), unavailable on this box"}
        try:
            import scipy.sparse as sp
            import scipy.sparse.linalg as spl
            c3s = host.gen_femlike(28, 0.58, 3)  # bounded sample: ILUTP fill makes the full 47^3 case take ~40 s of CPU
            A = sp.csr_matrix((c3s.values, c3s.colidx, c3s.rowptr), shape=(c3s.m, c3s.n)).tocsc()
            t0 = time.perf_counter()
            ilu = spl.spilu(A)  # defaults: drop_tol 1e-4, fill_factor 10 -- what ilu_set_default_options gives dgsisx
            t1 = time.perf_counter()
            ilu.solve(np.ones(c3s.m))
            t2 = time.perf_counter()
            t3 = time.perf_counter()
            _, _, lus = orc.ilu0(c3s)
            t4 = time.perf_counter()
            out["scipy_spilu_femlike28_fp64"] = {"factor_ms": (t1 - t0) * 1e3, "solve_ms": (t2 - t1) * 1e3, "rows": c3s.m, "stored": c3s.stored,
                                                 "cpu_ilu0_same_matrix_factor_ms": (t4 - t3) * 1e3,
                                                 "label": "NON-EQUIVALENT algorithm, bounded sample (FEM-like 28^3 instead of 47^3): "
                                                          "scipy.sparse.linalg.spilu = SuperLU gsitrf (ILUTP with pivoting, drop_tol 1e-4, "
                                                          "fill 10), the algorithm the reference's dgsisx runs"}
        except Exception as e:  # noqa: BLE001
            out["scipy_spilu_femlike28_fp64"] = {"unavailable": repr(e)}
    print(json.dumps(out))


def cpu_leg(which, cpus=None, omp_threads=None, timeout=240):
    env = dict(os.environ)
    if omp_threads:
        env["OMP_NUM_THREADS"] = str(omp_threads)
    else:
        env.pop("OMP_NUM_THREADS", None)

    def pre():
        if cpus:
            os.sched_setaffinity(0, cpus)
    try:
        r = subprocess.run([sys.executable, os.path.abspath(__file__), "--cpu-leg", which], env=env, preexec_fn=pre,
                           capture_output=True, text=True, timeout=timeout)
        return json.loads(r.stdout.strip().splitlines()[-1])
    except Exception as e:  # noqa: BLE001
        return {"failed": repr(e)}


# ------------------------------------------------------------------------------------ reference arm
def run_reference(args):
    """The reference's CPU SpMV path (test_spmv.c:168-180, mkl_sparse_d_mv under OpenMP) on this box's
    host cores.  MKL is not in the image and cannot be installed: the oracle's OpenMP restatement of
    that call is timed instead (kind "port").  Rank 0 only; the thread count is every core this process may run on
    (torchrun exports OMP_NUM_THREADS=1: the count is passed to the OpenMP region explicitly)."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    import orc
    from respasol_b200 import host
    cpus = all_host_cpus()
    threads = len(cpus)
    if args.gpus == 1:
        a = host.gen_stencil27(GRID, seed=2)
    else:
        a = host.gen_stencil27(GRID5, seed=2)
    x = orc.dlarnv(a.n) - 0.5
    bytes_per = spmv_bytes(a.stored, a.m, a.n, "d")
    W = max(args.warmup, 1)
    for _ in range(W):
        orc.spmv(a, x, "d", threads=threads)
    steps = max(1, min(args.steps, 200))  # bounded sample: each product is milliseconds on the host
    t0 = time.perf_counter()
    for _ in range(steps):
        y = orc.spmv(a, x, "d", threads=threads)
    dt = time.perf_counter() - t0
    val = bytes_per * steps / dt * 1e-9
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": "GB/s", "n_gpus": args.gpus, "steps": steps,
            "warmup": W, "ms_per_step": dt / steps * 1e3, "higher_is_better": True,
            "scaling": "weak" if args.gpus == 1 else "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": job_config(args.gpus),
            "cpu_baseline": {"value": val, "unit": "GB/s", "cores": threads, "kind": "port",
                             "sample": "%d products of the full matrix, OpenMP static rows on all %d host cores this process may use; "
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


def cold_rate(torch, h, make, x_of, precision, reps, target_bytes=400e6):
    """us per product with enough resident replicas rotated that no product finds its stream in L2"""
    a = make()
    from respasol_b200.gpu import DeviceCSR
    b = spmv_bytes(a.stored, a.m, a.n, precision)
    R = max(3, int(np.ceil(target_bytes / b)) + 1)
    xt = torch.float32 if precision == "s" else torch.float64
    mats = [DeviceCSR(a, precision) for _ in range(R)]
    xs = [torch.from_numpy(x_of(a)).cuda().to(xt) for _ in range(R)]
    ys = [torch.empty(a.m, dtype=xt, device="cuda") for _ in range(R)]
    for r in range(R):
        h.spmv(mats[r], xs[r], ys[r])
    t, _, _ = timed_loop(torch, lambda i: h.spmv(mats[i % R], xs[i % R], ys[i % R]), reps, lambda: None)
    return a, t / reps * 1e3, b, R, ys[0]


def per_call_cold(torch, fn, reps, flush):
    """the cuSPARSE leg's protocol, same for both sides: cudaEvents around EACH call, 5 warm-ups, a 256 MB memset
    between repetitions and outside the event pair (cold, partly dirty L2); median microseconds"""
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for _ in range(5):
        fn()
    ts = []
    for i in range(reps):
        flush.fill_(i & 0xff)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    return float(np.median(ts))


def cusparse_leg(torch, h, orc, host, reps=30):
    """cuSPARSE 12.x on the same box, same arrays, ONE protocol for both sides (per_call_cold).  The checker library
    oracle/libcusparse_ref.so restates the reference harness's call groups (GPU/spmv.cu:132-186, GPU/ilu0.cu:166-310)."""
    from respasol_b200.gpu import DeviceCSR
    out = {"protocol": "cudaEvents per call, 5 warm-ups, 256 MB memset between repetitions outside the event pair, median of %d; "
                       "cuSPARSE = best of ALG_DEFAULT / CSR_ALG1 / CSR_ALG2 after cusparseSpMV_preprocess" % reps}
    try:
        out["cusparse_version"] = int(orc.cusparse().cref_version())
        flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
        cases = (("c1_lap2d_1000", host.gen_laplace2d(1000)), ("c2_stencil27_100", host.gen_stencil27(GRID, seed=2)),
                 ("c4_powerlaw", host.gen_powerlaw(320000, 2.2, 64, 4, 100000, 4)))
        for name, a in cases:
            for prec in ("d", "s"):
                A = DeviceCSR(a, prec)
                x = torch.from_numpy(orc.dlarnv(a.n) - 0.5).cuda().to(A.values.dtype)
                y = torch.empty(a.m, dtype=x.dtype, device="cuda")
                ours = per_call_cold(torch, lambda: h.spmv(A, x, y), reps, flush)
                best, diff = None, 0.0
                for alg in (0, 1, 2):
                    y_ref, us = orc.cusparse_spmv(A, x, alg=alg, warm=5, reps=reps, flush=flush)
                    best = us[1] if best is None else min(best, us[1])
                    diff = max(diff, float((y - y_ref).abs().max() / y_ref.abs().max()))
                out["spmv_%s_%s" % (name, "fp64" if prec == "d" else "fp32")] = {
                    "ours_us": round(ours, 2), "cusparse_best_us": round(best, 2), "speedup": round(best / ours, 3), "rel_diff": diff}
                del A
        c3 = host.gen_femlike(47, 0.58, 3)
        for prec, dt in (("d", torch.float64), ("s", torch.float32)):
            A = DeviceCSR(c3, prec)
            pristine = A.values.clone()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

            def ev(fn):
                torch.cuda.synchronize()
                e0.record()
                fn()
                e1.record()
                torch.cuda.synchronize()
                return e0.elapsed_time(e1)
            b = torch.ones(c3.m, dtype=dt, device="cuda")
            z, yv = torch.empty_like(b), torch.empty_like(b)
            fa, so = [], []
            for _ in range(12):
                A.values.copy_(pristine)
                fa.append(ev(lambda: h.ilu0_factor(A, report_pivot=False)))
                so.append(ev(lambda: h.sptrsv_lu(A, A.values, b, z, yv)))
            cf = []
            for level in (0, 1):
                tl = []
                for _ in range(4):
                    v = pristine.clone()
                    _, _, _, ms_f = orc.cusparse_ilu0(A, v, level=level)
                    tl.append(ms_f)
                cf.append(float(np.median(tl)))
            _, y_ref, _, ms_s = orc.cusparse_spsv_lu(A, v, b, reps=10)
            out["ilu0_c3_%s" % ("fp64" if prec == "d" else "fp32")] = {
                "ours_factor_ms": float(np.median(fa[2:])), "cusparse_csrilu02_ms": min(cf),
                "ours_solve_LU_ms": float(np.median(so[2:])), "cusparse_spsv_LU_ms": float(ms_s),
                "speedup_factor": min(cf) / float(np.median(fa[2:])), "speedup_solve": float(ms_s) / float(np.median(so[2:])),
                "solve_rel_diff": float((yv - y_ref).abs().max() / y_ref.abs().max())}
            del A
        del flush
    except Exception as e:  # noqa: BLE001
        out["failed"] = repr(e)
    return out


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
    if args.graph and args.exchange != "nccl":
        raise SystemExit("--graph bakes the exchange epoch into the graph: NCCL exchange only (--exchange nccl)")

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

    def allmin(v):
        return -allmax(-v)

    h = Handle(local_rank)
    # every matrix of this bench is static while its products are timed (the reference harness's loop, spmv.cu:174-207,
    # is 50 products of one matrix): the promise RSG_OPT_SPMV_PDL = 2 lets a product's CTAs fetch their tiles while the
    # previous product drains.  --pdl 0 / 1 run the same workload without the promise / without dependent launches.
    h.set_option("SPMV_PDL", args.pdl)
    K, W = args.steps, max(args.warmup, 3)
    peak, peak_src = peaks()
    sampler = ClockSampler(local_rank)
    sampler.start()
    extra = {}

    def dist_ops(a, starts, xs_local, replicas=REPLICAS, overlap=False, chain=False):
        """`replicas` row-block operators of this rank's block `a` (rotated by the timed loop)"""
        ops = []
        for _ in range(replicas):
            if world > 1 and chain:
                op = RowBlockSpmv(a, starts, "p2p-chain", lambda c: DeviceCSR(c, "d", dev), lambda A, x, y: h.spmv(A, x, y), dev,
                                  torch.float64, push=P2P_FNS, side=side)
                ops.append(op)
                continue
            if world > 1:
                op = RowBlockSpmv(a, starts, "p2p" if args.exchange.startswith("p2p") else "halo", lambda c: DeviceCSR(c, "d", dev),
                                  lambda A, x, y: h.spmv(A, x, y), dev, torch.float64,
                                  view=(lambda A, r0, r1: A.view_rows(r0, r1)) if overlap else None, push=P2P_FNS, side=side,
                                  proto=args.proto)
                if op.p2p is not None and op.p2p[1] is not None:
                    h.set_option("SPMV_HALO_START", op.plan.halo_base)  # marks the tiles that must wait for the peers
            else:
                op = RowBlockSpmv(a, starts, "halo", lambda c: DeviceCSR(c, "d", dev), lambda A, x, y: h.spmv(A, x, y), dev, torch.float64)
            op.set_x(torch.from_numpy(np.ascontiguousarray(xs_local)).to(dev))
            if getattr(op, "cdist", None) is None:  # (rsg_dist_create has built the plan of a fused operator already)
                for A in ([piece for piece, _, _ in op.split] if op.split else [op.A]):
                    h.spmv_analysis(A)
            ops.append(op)
        h.set_option("SPMV_HALO_START", 0)
        return ops

    def full_parity(a, x_global, ops, what):
        """EVERY row of this rank's block against the oracle (which reads the global x): bit-exact where the kernel sums a
        row left to right like the oracle (the window kernel's thread-per-row path), 1e-12 otherwise"""
        ref = orc.spmv(a, x_global, "d", threads=len(all_host_cpus()))
        exact = True
        for op in ops:
            got = op.y.cpu().numpy()
            assert np.abs(got - ref).max() <= 1e-12 * np.abs(ref).max(), "rank %d: %s differs from the oracle" % (rank, what)
            exact = exact and bool(np.array_equal(got, ref))
        return exact

    if world > 1:
        hp = Handle(local_rank)  # the push kernels of the non-fused modes run on a stream of their own, beside the product
        side = torch.cuda.Stream(device=dev)
        with torch.cuda.stream(side):
            hp.use_current_stream()
        if args.exchange == "p2p-barrier":
            hp.use_current_stream()
        push = hp.halo_push
        P2P_FNS = {"p2p": (push, h.spmv_halo_wait, h.spmv_halo_fuse), "p2p-side": (push, h.spmv_halo_wait),
                   "p2p-barrier": (push, None), "nccl": (push, None)}[args.exchange]
    else:
        hp, side, P2P_FNS = None, None, None

    # ---- the timed workload
    if world == 1:
        a = host.gen_stencil27(GRID, seed=2)
        xg = orc.dlarnv(a.n) - 0.5
        bytes_total = float(spmv_bytes(a.stored, a.m, a.n, "d"))
        mats = [DeviceCSR(a, "d", dev) for _ in range(REPLICAS)]
        xd = [torch.from_numpy(xg).to(dev) for _ in range(REPLICAS)]
        yd = [torch.empty(a.m, dtype=torch.float64, device=dev) for _ in range(REPLICAS)]

        def step(i):
            r = i % REPLICAS
            h.spmv(mats[r], xd[r], yd[r])
        for r in range(REPLICAS):
            h.spmv_analysis(mats[r])
        launches_per_step = 1
        scaling = "weak"
        stored_total = a.stored
    else:
        cuts = [(r * GRID5 + world // 2) // world for r in range(world + 1)]
        a = host.gen_stencil27(GRID5, GRID5, GRID5, cuts[rank], cuts[rank + 1], seed=2)
        starts5 = np.array(cuts, dtype=np.int64) * GRID5 * GRID5
        xg = orc.dlarnv(GRID5 ** 3) - 0.5
        chained = args.mode == "chain" and args.exchange == "p2p"
        ops = dist_ops(a, starts5, xg[starts5[rank]:starts5[rank + 1]], overlap=args.overlap, chain=chained)
        stored_total = allsum(float(a.stored))
        bytes_total = stored_total * 12 + (GRID5 ** 3 + 1) * 4 + 2 * 8 * GRID5 ** 3
        if chained:
            for op in ops:
                op.set_x_global(xg)
        if args.graph:
            for op in ops:
                for _ in range(3):  # eager steps first: plans, NCCL channels, allocator all settled
                    op.step()
                op.capture(h.use_current_stream)
            extra["cuda_graph"] = "one graph per step (pack + NCCL all-gather + SpMV launches), replayed"

        def step(i):
            if chained:
                ops[i % REPLICAS].step_chain(CHAIN_ALPHA)
            else:
                ops[i % REPLICAS].step()
        launches_per_step = len(ops[0].split) if ops[0].split else 1
        launches_per_step += 1 if (ops[0].p2p is not None and len(ops[0].p2p) < 3) else 0
        scaling = "strong"
        extra["exchange"] = ops[0].describe()

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
    ms = allmax(ms)
    value = bytes_total * K / (ms * 1e-3) * 1e-9
    # keep the load up a little longer so the 20 ms clock poll sees the kernel even for a short K: a FIXED number of
    # steps, identical on every rank (a clock-bounded loop would let ranks disagree by a batch and drift apart)
    extra_steps = int(min(20000, max(0.0, 0.4 - (w1 - w0)) / max(ms / K * 1e-3, 1e-6)))
    if world > 1:
        t = torch.tensor([extra_steps], dtype=torch.int64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        extra_steps = int(t.item())
    for i in range(extra_steps):
        step(K + i)
    torch.cuda.synchronize()
    w1b = time.time()
    barrier()
    c0 = time.perf_counter()  # what the host needs to ISSUE a step: a short burst into an empty launch queue
    for i in range(64):
        step(K + extra_steps + i)
    host_issue_us = (time.perf_counter() - c0) / 64 * 1e6
    torch.cuda.synchronize()
    barrier()
    clocks = sampler.summary(w0, max(w1, w1b))
    extra["host_issue_us_per_step"] = round(host_issue_us, 2)
    h.synchronize()  # raises RSG_DEPENDENCY_TIMEOUT if a halo wait of the timed steps gave up

    # ---- full-vector parity of what was just timed
    if world == 1:
        ref = orc.spmv(a, xg, "d", threads=len(all_host_cpus()))
        exact = True
        for yv in yd:
            got = yv.cpu().numpy()
            assert np.abs(got - ref).max() <= 1e-12 * np.abs(ref).max(), "bench result differs from the oracle"
            exact = exact and bool(np.array_equal(got, ref))
        extra["parity_full"] = True
        extra["parity_bit_exact"] = exact
    elif chained:
        # restart both chains from the same x and walk CHAIN_PARITY_STEPS steps next to the oracle's chain on the FULL matrix
        a_full = host.gen_stencil27(GRID5, seed=2)
        ref = xg.copy()
        nthr = max(1, len(all_host_cpus()) // world)
        for _ in range(CHAIN_PARITY_STEPS):
            ref = orc.spmv(a_full, ref, "d", alpha=CHAIN_ALPHA, threads=nthr)
        del a_full
        exact = True
        for op in ops:
            op.set_x_global(xg)
            for _ in range(CHAIN_PARITY_STEPS):
                op.step_chain(CHAIN_ALPHA)
            h.synchronize()
            got = op.x_local().cpu().numpy()
            want = ref[starts5[rank]:starts5[rank + 1]]
            assert np.abs(got - want).max() <= 1e-12 * np.abs(want).max(), "rank %d: the chained configs[4] products differ from the oracle" % rank
            exact = exact and bool(np.array_equal(got, want))
        extra["parity_full"] = bool(allmin(1.0) > 0)
        extra["parity_bit_exact"] = bool(allmin(1.0 if exact else 0.0) > 0)
        extra["parity"] = ("every entry of x after %d chained products (all %d rows) against the oracle's chain on the global matrix; the "
                           "halo of every step is what the neighbours' tiles pushed" % (CHAIN_PARITY_STEPS, GRID5 ** 3))
        # the same matrix with SEPARATE products (static x; the push precedes the product inside its launch): round 1's protocol
        ops_s = dist_ops(a, starts5, xg[starts5[rank]:starts5[rank + 1]])
        for i in range(6):
            ops_s[i % REPLICAS].step()
        Ks = max(50, min(K, 1000))
        mss, _, _ = timed_loop(torch, lambda i: ops_s[i % REPLICAS].step(), Ks, barrier)
        mss = allmax(mss)
        exs = full_parity(a, xg, ops_s, "the distributed configs[4] product")
        extra["c5_strong_separate_products_fp64"] = {"us_per_product": mss / Ks * 1e3, "GBps_aggregate": bytes_total / (mss / Ks * 1e-3) * 1e-9,
                                                      "parity_full": True, "parity_bit_exact": bool(allmin(1.0 if exs else 0.0) > 0),
                                                      "exchange": ops_s[0].describe()["mode"]}
        del ops_s
    else:
        exact = full_parity(a, xg, ops, "the distributed configs[4] product")
        extra["parity_full"] = bool(allmin(1.0) > 0)
        extra["parity_bit_exact"] = bool(allmin(1.0 if exact else 0.0) > 0)
        extra["parity"] = "every row of every rank's block (all %d rows of configs[4]) against the oracle on the global x" % GRID5 ** 3

    # ---- roofline of the dominant kernel
    per_launch_ms = ms / K
    algo_used = h.info("SPMV_ALGO_USED")
    traffic = None
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tp):
        with open(tp) as f:
            traffic = json.load(f).get("spmv_window_fp64_c2_bytes_per_launch" if world == 1 else None)
    my_bytes = float(spmv_bytes(a.stored, a.m, a.m if world > 1 else a.n, "d"))
    achieved = my_bytes / (per_launch_ms * 1e-3) * 1e-9  # this GPU's launch: its rows, its slice of x
    kname = {8: "spmv_window_kernel<double,double,%s,4,%s>" % ("64" if world > 4 else "128", "true" if world > 4 else "false"), 7: "spmv_tile_kernel<double,double,128,2304,512,7,%s>" % ("true" if world > 1 else "false"),
             3: "spmv_vector_kernel<double,double>"}.get(algo_used, "algo %d" % algo_used)
    roofline = {"bound": "hbm", "kernel": kname, "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": traffic, "peak_source": peak_src, "algorithmic_bytes_per_launch": my_bytes,
                "frac_of_8TBs_nominal": achieved / 8000.0, "us_per_launch": per_launch_ms * 1e3,
                "dram_GBps": (traffic / (per_launch_ms * 1e-3) * 1e-9) if traffic else None,
                "note": "achieved = ALGORITHMIC CSR bytes / time (the contract).  The window kernel moves fewer bytes than that: "
                        "the plan stores a 16-bit window-local column per entry (2 B instead of colidx's 4 B) and the x window of a "
                        "tile travels once per tile; `traffic` (ncu dram__bytes read+write per launch) / time = dram_GBps is what HBM "
                        "really sustained, and it is that figure that sits at the measured copy peak"}

    e2e = None
    cpu_baseline = None
    if world > 1:
        # ---- round 1's curve, kept: weak scaling of configs[1] per GPU (every rank owns a 100x100x100 slab of a 100x100x(100 N) grid)
        nz = GRID * world
        aw = host.gen_stencil27(GRID, GRID, nz, rank * GRID, (rank + 1) * GRID, seed=2)
        xw = orc.dlarnv(GRID * GRID * nz) - 0.5
        opsw = dist_ops(aw, np.arange(world + 1, dtype=np.int64) * aw.m, xw[rank * aw.m:(rank + 1) * aw.m])
        for i in range(6):
            opsw[i % REPLICAS].step()
        Kw = max(50, min(K, 1000))
        msw, _, _ = timed_loop(torch, lambda i: opsw[i % REPLICAS].step(), Kw, barrier)
        msw = allmax(msw)
        bw = allsum(float(spmv_bytes(aw.stored, aw.m, aw.m, "d")))
        exw = full_parity(aw, xw, opsw, "the distributed configs[1]-per-GPU product")
        extra["c2_weak_fp64"] = {"workload": "27-point 100x100x%d, one 100^3 slab (configs[1]) per GPU" % nz, "us_per_product": msw / Kw * 1e3,
                                 "GBps_aggregate": bw / (msw / Kw * 1e-3) * 1e-9, "scaling": "weak", "parity_full": True,
                                 "parity_bit_exact": bool(allmin(1.0 if exw else 0.0) > 0)}
        del opsw, aw
    full = a if world == 1 else host.gen_stencil27(GRID, seed=2)
    xfull = xg if world == 1 else orc.dlarnv(full.n) - 0.5
    if world == 1:
        # ---- other precisions / configs (single-GPU kernels; reported, not the headline)
        reps = max(20, min(K, 500))
        extra["spmv_fp64"] = {"GBps": achieved, "us": per_launch_ms * 1e3, "GFLOPs": 2.0 * a.stored / (per_launch_ms * 1e-3) * 1e-9,
                              "frac_of_measured_peak": achieved / peak, "algo": algo_used}
        ab = {}
        for mode in (1, 0, 2):  # plain launches / dependent launch, plan data early / + values early
            h.set_option("SPMV_PDL", mode)
            for i in range(4):
                step(i)
            t, _, _ = timed_loop(torch, step, reps, lambda: None)
            ab["pdl_%d_us" % mode] = t / reps * 1e3
        h.set_option("SPMV_PDL", args.pdl)
        extra["spmv_fp64"]["launch_modes_us"] = ab
        for p in ("s", "m"):
            _, us, b, R, _ = cold_rate(torch, h, lambda: full, lambda m_: xfull, p, reps)
            extra["spmv_" + {"s": "fp32", "m": "mixed"}[p]] = {"GBps": b / us * 1e-3, "us": us, "GFLOPs": 2.0 * full.stored / us * 1e-3,
                                                                "frac_of_measured_peak": b / us * 1e-3 / peak, "replicas": R,
                                                                "algo": h.info("SPMV_ALGO_USED")}
        # configs[4] on ONE GPU (the N = 1 point of the strong-scaling curve the N > 1 lines report)
        if not args.no_c5:
            a5, us5, b5, R5, y5 = cold_rate(torch, h, lambda: host.gen_stencil27(GRID5, seed=2), lambda m_: orc.dlarnv(m_.n) - 0.5, "d",
                                            max(20, min(K, 200)), target_bytes=0)
            ref5 = orc.spmv(a5, orc.dlarnv(a5.n) - 0.5, "d", threads=len(all_host_cpus()))
            got5 = y5.cpu().numpy()
            assert np.abs(got5 - ref5).max() <= 1e-12 * np.abs(ref5).max(), "configs[4] result differs from the oracle"
            extra["c5_strong_fp64"] = {"workload": "27-point 134^3, %d entries, 1 GPU" % a5.stored, "us_per_product": us5,
                                       "GBps_aggregate": b5 / us5 * 1e-3, "frac_of_measured_peak": b5 / us5 * 1e-3 / peak,
                                       "GFLOPs": 2.0 * a5.stored / us5 * 1e-3, "scaling": "strong", "parity_full": True,
                                       "parity_bit_exact": bool(np.array_equal(got5, ref5)), "replicas": R5}
            del a5, y5, ref5, got5
            torch.cuda.empty_cache()
        # configs[0] and configs[3], COLD L2 (replicas rotated: > 400 MB of distinct streams) and warm (one resident copy)
        for key, make, xof in (("spmv_c1_lap2d_fp64", lambda: host.gen_laplace2d(1000), lambda m_: np.ones(m_.n)),
                               ("spmv_c4_powerlaw_fp64", lambda: host.gen_powerlaw(320000, 2.2, 64, 4, 100000, 4), lambda m_: orc.dlarnv(m_.n))):
            am, us_cold, b, R, y0 = cold_rate(torch, h, make, xof, "d", reps)
            algo_m = h.info("SPMV_ALGO_USED")
            refm = orc.spmv(am, xof(am), "d", threads=len(all_host_cpus()))
            assert np.abs(y0.cpu().numpy() - refm).max() <= 1e-12 * np.abs(refm).max(), key + " differs from the oracle"
            A1 = DeviceCSR(am, "d", dev)
            x1 = torch.from_numpy(xof(am)).to(dev)
            y1 = torch.empty(am.m, dtype=torch.float64, device=dev)
            for _ in range(3):
                h.spmv(A1, x1, y1)
            tw, _, _ = timed_loop(torch, lambda i: h.spmv(A1, x1, y1), reps, lambda: None)
            extra[key] = {"us_cold": us_cold, "GBps_cold": b / us_cold * 1e-3, "frac": b / us_cold * 1e-3 / peak, "replicas": R,
                          "us_L2_warm": tw / reps * 1e3, "GBps_L2_warm": b / (tw / reps * 1e3) * 1e-3, "algo": algo_m,
                          "stored": am.stored, "max_row": h.info("SPMV_MAX_ROW"), "parity_full": True}
            if key.startswith("spmv_c4"):
                for alg in ("rowblock", "merge"):
                    h.set_option("SPMV_ALGO", ALGO[alg])
                    for _ in range(2):
                        h.spmv(A1, x1, y1)
                    t4, _, _ = timed_loop(torch, lambda i: h.spmv(A1, x1, y1), max(5, reps // 10), lambda: None)
                    h.set_option("SPMV_ALGO", 0)
                    extra[key]["us_L2_warm_" + alg] = t4 / max(5, reps // 10) * 1e3
            del A1, am
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
        # thin levels: ILU0 + solves on the 2-D 5-point 1000^2 -- 1999 wavefronts
        c1 = host.gen_laplace2d(1000)
        A1 = DeviceCSR(c1, "d", dev)
        pristine = A1.values.clone()
        b1 = torch.ones(c1.m, dtype=torch.float64, device=dev)
        z1, y1 = torch.empty_like(b1), torch.empty_like(b1)
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
        tf, ts = [], []
        for it in range(6):
            A1.values.copy_(pristine)
            torch.cuda.synchronize()
            ev[0].record(); h.ilu0_factor(A1, report_pivot=False); ev[1].record(); torch.cuda.synchronize()
            tf.append(ev[0].elapsed_time(ev[1]))
            ev[0].record(); h.sptrsv_lu(A1, A1.values, b1, z1, y1); ev[1].record(); torch.cuda.synchronize()
            ts.append(ev[0].elapsed_time(ev[1]))
        extra["ilu0_thin_levels_lap2d_1000_fp64"] = {"factor_ms": float(np.median(tf[1:])), "solve_LU_ms": float(np.median(ts[1:])),
                                                     "levels_L": h.info("LEVELS_L"), "n": c1.m, "stored": c1.stored}
        del A1, c1
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
        if not args.no_cusparse:
            extra["cusparse"] = cusparse_leg(torch, h, orc, host)
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

    def copy_us(dst, src, reps=20, concurrent=False):
        for _ in range(3):
            dst.copy_(src, non_blocking=True)
        torch.cuda.synchronize()
        if concurrent:
            barrier()
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
    n_spin = int(0.3 / 50e-6)
    for _ in range(n_spin):
        h.spmv(spin_A, spin_x, spin_y)
    torch.cuda.synchronize()
    pcie = {}
    if world > 1:  # one rank at a time, then all ranks at once: is the host side (root complex / memory) the shared limit?
        for r in range(world):
            barrier()
            if r == rank:
                pcie["h2d_alone_us"] = round(copy_us(spin_x, xh_t), 1)
                pcie["d2h_alone_us"] = round(copy_us(yh_t, spin_y), 1)
        barrier()
        pcie["h2d_all_ranks_at_once_us"] = round(allmax(copy_us(spin_x, xh_t, concurrent=True)), 1)
        pcie["d2h_all_ranks_at_once_us"] = round(allmax(copy_us(yh_t, spin_y, concurrent=True)), 1)
        pcie["h2d_alone_us_max_over_ranks"] = round(allmax(pcie["h2d_alone_us"]), 1)
        pcie["d2h_alone_us_max_over_ranks"] = round(allmax(pcie["d2h_alone_us"]), 1)
    else:
        pcie["h2d_alone_us"] = round(copy_us(spin_x, xh_t), 1)
        pcie["d2h_alone_us"] = round(copy_us(yh_t, spin_y), 1)
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
    refy = orc.spmv(full64, xfull, "d", threads=len(all_host_cpus()))
    assert np.abs(yh - refy).max() <= 1e-12 * np.abs(refy).max()
    pcie["link_during_rounds"] = sorted(set(links))
    pcie["note"] = "plain cudaMemcpyAsync of the same 8 MB buffers on this box, measured just before the rounds"
    e2e = {"value": world * spmv_bytes(full64.stored, full64.m, full64.n, "d") / edt * 1e-9, "unit": "GB/s",
           "h2d_bytes_per_step": 8 * full64.n * world, "d2h_bytes_per_step": 8 * full64.m * world, "ms_per_step": edt * 1e3,
           "host_cpus": ("%d GPU-local of %d" % (len(near), len(all_cpus))) if near else "unbound",
           "ms_per_step_rounds": [round(r * 1e3, 4) for r in rounds], "pcie": pcie, "parity_full": True,
           "workload": "configs[1] (27-point 100^3) on every rank" if world > 1 else "configs[1]",
           "note": "rsg_matrix_spmv_host on every rank: matrix resident in HBM (uploaded once, as GPU/spmv.cu:111-118 does), "
                   "x H2D (from pinned host memory) + SpMV + y D2H (to pinned host memory) inside every timed step"}
    if world == 1 and not args.no_cpu:
        # ---- CPU baseline on this box's host cores (bounded samples): all cores, then the reference's own protocol
        cpus = all_host_cpus()
        threads = len(cpus)
        n_cpu, cdt, _ = cpu_spmv_rate(orc, full64, xfull, "d", threads, 8.0, 200)
        cpu_baseline = {"value": spmv_bytes(full64.stored, full64.m, full64.n, "d") / cdt * 1e-9, "unit": "GB/s",
                        "cores": threads, "kind": "port", "ms_per_product": cdt * 1e3,
                        "sample": "%d fp64 products of the full config-2 matrix (~8 s), OpenMP static rows on all %d host cores; "
                                  "stands in for mkl_sparse_d_mv (test_spmv.c:168-170), MKL absent" % (n_cpu, threads)}
        n32, c32, _ = cpu_spmv_rate(orc, as_f32(full64), xfull.astype(np.float32), "s", threads, 4.0, 200)
        cpu_baseline["fp32"] = {"value": spmv_bytes(full64.stored, full64.m, full64.n, "s") / c32 * 1e-9, "unit": "GB/s",
                                "ms_per_product": c32 * 1e3, "sample": "%d fp32 products, same threads (test_spmv.c:177-183)" % n32}
        cpu_baseline["reference_protocol_4_pinned_threads"] = cpu_leg("pinned4", cpus=set(cpus[:4]), omp_threads=4)
        ilu = cpu_leg("ilu", cpus=set(cpus[:1]), omp_threads=1)
        cpu_baseline.update(ilu if isinstance(ilu, dict) else {})
        if "cpu_ilu0_c3_fp64" in cpu_baseline:
            extra["cpu_ilu0_c3_fp64_ms"] = cpu_baseline["cpu_ilu0_c3_fp64"].get("factor_ms")
            extra["cpu_sptrsv_LU_c3_fp64_ms"] = cpu_baseline["cpu_ilu0_c3_fp64"].get("solve_LU_ms")
    sampler.stop_flag = True
    barrier()
    if rank == 0:
        cfg = job_config(world)
        line = {"metric": METRIC, "value": value, "unit": "GB/s", "n_gpus": world, "steps": K, "warmup": W,
                "ms_per_step": ms / K, "higher_is_better": True, "scaling": scaling, "vs_baseline": None, "dtype": "f64",
                "data": "synthetic", "config": cfg,
                "gflops": 2.0 * stored_total * K / (ms * 1e-3) * 1e-9,
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
    ap.add_argument("--no-cpu", action="store_true", help="skip the host-CPU baseline legs")
    ap.add_argument("--no-c5", action="store_true", help="N = 1: skip the 64M-entry matrix (BASELINE configs[4]) leg")
    ap.add_argument("--no-cusparse", action="store_true", help="N = 1: skip the cuSPARSE comparison leg")
    ap.add_argument("--graph", action="store_true", help="N>1: replay each step from a CUDA graph (NCCL exchange only)")
    ap.add_argument("--overlap", action="store_true", help="N>1: exchange on a side stream while the interior rows are multiplied")
    ap.add_argument("--exchange", default="p2p", choices=["p2p", "p2p-side", "p2p-barrier", "nccl"], help="N>1: x-halo exchange: peer-memory push or NCCL all-gather")
    ap.add_argument("--pdl", type=int, default=2, choices=[0, 1, 2], help="RSG_OPT_SPMV_PDL for the timed products (2 = matrix values promised static)")
    ap.add_argument("--proto", default="auto", choices=["auto", "pull", "packets", "packets-fused", "flags"],
                    help="N>1, fused exchange: pull = the product's first CTAs fetch the halo (no remote store in the product); "
                         "packets = {value, step} packets pushed and unpacked by the halo tiles; flags = push + fence + arrival flags")
    ap.add_argument("--mode", default="separate", choices=["chain", "separate"],
                    help="N>1: separate = independent products of a static x, the push at the head of each launch (default; round 1's "
                         "protocol); chain = products x <- A x / 52 whose boundary rows are pushed by the tiles that compute them "
                         "(measured no faster: DESIGN.md 6)")
    ap.add_argument("--cpu-leg", default=None, help=argparse.SUPPRESS)
    args = ap.parse_args()
    if args.cpu_leg:
        cpu_leg_child(args.cpu_leg)
    elif args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
