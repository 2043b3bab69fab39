#!/usr/bin/env python
"""Competitor timing on the same B200: the reference harness's cuSPARSE calls (oracle/libcusparse_ref.so)
next to libreaspasol_gpu on the same device arrays.  MEASUREMENT INFRASTRUCTURE (lives in tests/ because it
loads oracle/); not collected by pytest.
    python tests/compare_cusparse.py [--reps 50] [--out gpurun_out/cusparse_compare.json]
Protocol: cudaEvents around each call, 5 warm-ups, a 256 MB memset between repetitions and
outside the event pair so every product starts with a cold L2; median / min / mean of `reps`."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402
import torch  # noqa: E402
import orc  # noqa: E402
from respasol_b200 import host  # noqa: E402
from respasol_b200.gpu import DeviceCSR, Handle  # noqa: E402


def time_ours(fn, reps, flush):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for _ in range(5):
        fn()
    ts = []
    for i in range(reps):
        if flush is not None:
            flush.fill_(i & 0xff)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    ts = np.array(ts)
    return [float(ts.mean()), float(np.median(ts)), float(ts.min())]


def spmv_case(h, name, a, reps, flush, out):
    for prec in ("d", "s"):
        A = DeviceCSR(a, prec)
        x = torch.from_numpy(orc.dlarnv(a.n) - 0.5).cuda().to(A.values.dtype)
        y = torch.empty(a.m, dtype=x.dtype, device="cuda")
        h.spmv(A, x, y)
        ours = time_ours(lambda: h.spmv(A, x, y), reps, flush)
        row = {"bytes": A.bytes(), "stored": a.stored, "rows": a.m, "ours_us_mean_med_min": ours,
               "ours_algo": h.info("SPMV_ALGO_USED")}
        best = None
        for alg, an in ((0, "default"), (1, "csr_alg1"), (2, "csr_alg2")):
            y_ref, us = orc.cusparse_spmv(A, x, alg=alg, warm=5, reps=reps, flush=flush)
            row["cusparse_%s_us_mean_med_min" % an] = us
            err = float((y - y_ref).abs().max() / y_ref.abs().max())
            row["rel_diff_vs_%s" % an] = err
            best = us[1] if best is None else min(best, us[1])
        if prec == "d":  # the transposed product: ours (cached transposed pattern + value gather, deterministic) vs op = T
            xt_ = torch.from_numpy(orc.dlarnv(a.m) - 0.5).cuda()
            yt_ = torch.empty(a.n, dtype=torch.float64, device="cuda")
            h.spmv_t(A, xt_, yt_)
            row["transposed_ours_us_mean_med_min"] = time_ours(lambda: h.spmv_t(A, xt_, yt_), reps, flush)
            h.set_option("TRANSPOSE_STATIC", 1)  # the caller's promise that the values stay put: no per-call gather
            h.spmv_t(A, xt_, yt_)
            row["transposed_static_ours_us_mean_med_min"] = time_ours(lambda: h.spmv_t(A, xt_, yt_), reps, flush)
            h.set_option("TRANSPOSE_STATIC", 0)
            yr_, ust = orc.cusparse_spmv(A, xt_, warm=5, reps=reps, flush=flush, transpose=True)
            row["transposed_cusparse_us_mean_med_min"] = ust
            row["transposed_rel_diff"] = float((yt_ - yr_).abs().max() / yr_.abs().max())
        row["ours_GBps"] = A.bytes() / ours[1] / 1e3
        row["cusparse_best_GBps"] = A.bytes() / best / 1e3
        row["speedup_vs_best_cusparse"] = best / ours[1]
        out["spmv_%s_%s" % (name, "fp64" if prec == "d" else "fp32")] = row
        print(name, prec, json.dumps(row), flush=True)
        del A


def ilu_case(h, name, a, reps, out):
    for prec, dt in (("d", torch.float64), ("s", torch.float32)):
        A = DeviceCSR(a, prec)
        pristine = A.values.clone()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

        def ev(fn):
            torch.cuda.synchronize()
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize()
            return e0.elapsed_time(e1)
        an, fa, so = [], [], []
        b = torch.ones(a.m, dtype=dt, device="cuda")
        z, y = torch.empty_like(b), torch.empty_like(b)
        for i in range(reps + 2):
            A.values.copy_(pristine)
            h.lib.rsg_forget_matrix(h.h, A.rowptr.data_ptr())
            t_an = ev(lambda: (h.ilu0_analysis(A), h.sptrsv_analysis(A)))
            t_fa = ev(lambda: h.ilu0_factor(A, report_pivot=False))
            t_so = ev(lambda: h.sptrsv_lu(A, A.values, b, z, y))
            if i >= 2:
                an.append(t_an); fa.append(t_fa); so.append(t_so)
        row = {"rows": a.m, "stored": a.stored, "levels": h.info("LEVELS_L"),
               "ours_ms": {"analysis_all": float(np.median(an)), "factor": float(np.median(fa)), "solve_LU": float(np.median(so))}}
        ours_lu = A.values.clone()
        for level, ln in ((0, "no_level"), (1, "use_level")):
            ta, tf = [], []
            for i in range(max(3, reps // 3)):
                v = pristine.clone()
                _, _, ms_a, ms_f = orc.cusparse_ilu0(A, v, level=level)
                ta.append(ms_a); tf.append(ms_f)
            row["cusparse_csrilu02_%s_ms" % ln] = {"analysis": float(np.median(ta)), "factor": float(np.median(tf))}
        z_ref, y_ref, ms_a, ms_s = orc.cusparse_spsv_lu(A, v, b, reps=reps)
        row["cusparse_spsv_ms"] = {"analysis_LU": ms_a, "solve_LU": ms_s}
        row["factor_rel_diff"] = float((ours_lu - v).abs().max() / v.abs().max())
        row["solve_rel_diff"] = float((y - y_ref).abs().max() / y_ref.abs().max())
        row["speedup_factor_vs_best_csrilu02"] = min(row["cusparse_csrilu02_no_level_ms"]["factor"],
                                                     row["cusparse_csrilu02_use_level_ms"]["factor"]) / row["ours_ms"]["factor"]
        row["speedup_solve_vs_spsv"] = ms_s / row["ours_ms"]["solve_LU"]
        out["ilu0_%s_%s" % (name, "fp64" if prec == "d" else "fp32")] = row
        print(name, prec, json.dumps(row), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reps", type=int, default=50)
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "cusparse_compare.json"))
    ap.add_argument("--skip-c5", action="store_true")
    args = ap.parse_args()
    h = Handle(0)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    out = {"cusparse_version": orc.cusparse().cref_version(), "device": torch.cuda.get_device_name(0),
           "protocol": "cudaEvents per call, 5 warm-ups, 256 MB memset between repetitions (cold L2), [mean, median, min] us"}
    spmv_case(h, "c1_lap2d_1000", host.gen_laplace2d(1000), args.reps, flush, out)
    spmv_case(h, "c2_stencil27_100", host.gen_stencil27(100, seed=2), args.reps, flush, out)
    spmv_case(h, "c4_powerlaw", host.gen_powerlaw(320000, 2.2, 64, 4, 100000, 4), args.reps, flush, out)
    if not args.skip_c5:
        spmv_case(h, "c5_stencil27_134", host.gen_stencil27(134, seed=2), max(10, args.reps // 2), flush, out)
    ilu_case(h, "c3_femlike_47", host.gen_femlike(47, 0.58, 3), 10, out)
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    with open(args.out, "w") as f:
        json.dump(out, f, indent=1)
    h.close()


if __name__ == "__main__":
    main()
