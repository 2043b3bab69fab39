#!/usr/bin/env python
"""Summarise an .ncu-rep (read on the CPU box): key raw metrics + top stall instructions.
    python tools/ncu_summary.py gpurun_out/prof.ncu-rep [n_top]"""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
ntop = int(sys.argv[2]) if len(sys.argv) > 2 else 25
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, data = rows[0], rows[1], rows[2:]
want = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers", "launch__grid_size",
        "launch__block_size", "launch__shared_mem_per_block_dynamic", "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__inst_executed.sum", "sm__cycles_elapsed.max",
        "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum",
        "smsp__average_warp_latency_per_inst_issued.ratio", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "lts__t_sectors_srcunit_tex_op_read.sum", "lts__t_bytes.sum",
        "l1tex__lsu_writeback_active.avg.pct_of_peak_sustained_elapsed", "l1tex__m_xbar2l1tex_read_bytes.sum"]
for i, h in enumerate(hdr):
    if h in want or h.startswith("smsp__average_warps_issue_stalled") and h.endswith("per_issue_active.ratio"):
        vals = [d[i] for d in data]
        if h.startswith("smsp__average_warps_issue_stalled"):
            try:
                if max(float(v) for v in vals) < 0.3:
                    continue
            except ValueError:
                pass
        print("%-75s %-8s %s" % (h[:75], units[i], [v[:60] for v in vals]))
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hdr = rows[1]
ix = {n: hdr.index(n) for n in ["Source", "# Samples", "stall_long_sb", "stall_barrier", "stall_wait", "stall_short_sb", "stall_lg", "stall_mio", "Instructions Executed"]}
out = []
for r in rows[2:]:
    if r and r[0] == "Kernel Name":
        break
    if len(r) > ix["# Samples"]:
        try:
            out.append((int(r[ix["# Samples"]]), len(out), r[ix["Source"]].strip()[:60], {k: r[v] for k, v in ix.items() if k.startswith("stall") and r[v] not in ("0", "")}, r[ix["Instructions Executed"]]))
        except ValueError:
            pass
tot = sum(o[0] for o in out)
print("total samples", tot, "instructions", len(out))
for o in sorted(out, reverse=True)[:ntop]:
    print("%5d (%4.1f%%) #%-5d %-60s %s exec=%s" % (o[0], 100.0 * o[0] / max(tot, 1), o[1], o[2], o[3], o[4]))
