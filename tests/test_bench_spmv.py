"""bench_spmv: the reference's 36-matrix registered-benchmark sweep (bench_spmv.cc, bench_consts.h) with
google-benchmark-compatible output.  The CPU-path checks run anywhere; the GPU run is marked."""
import json
import os
import subprocess

import numpy as np
import pytest

from respasol_b200 import host

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "respasol_b200", "bin", "bench_spmv")


def run(*args):
    return subprocess.run([EXE, *args], capture_output=True, text=True, timeout=600)


def test_lists_36_matrices_per_precision():
    r = run("--benchmark_list_tests")
    names = r.stdout.split()
    assert r.returncode == 0 and len(names) == 72
    assert names[0] == "BMMatrixDouble/0" and names[35] == "BMMatrixDouble/35" and names[-1] == "BMMatrixSingle/35"
    assert "BMMatrixDouble/36" not in names  # the reference's DenseRange(0, 36) runs one past its table


def check_schema(doc, expect):
    assert set(doc) == {"context", "benchmarks"}
    assert [b["name"] for b in doc["benchmarks"]] == expect
    for b in doc["benchmarks"]:
        assert b["run_type"] == "iteration" and b["time_unit"] == "ms" and b["iterations"] >= 1
        assert b["real_time"] > 0 and b["cpu_time"] > 0 and b["rows"] > 0 and b["stored"] >= b["rows"]
        assert b["source"] == "synthetic"


def test_cpu_path_json_schema(tmp_path):
    out = tmp_path / "o.json"
    r = run("--cpu", "--scale=0.01", "--iterations=2", "--warmup=1", "--benchmark_filter=/(1|8|30)$", "--benchmark_out=%s" % out,
            "--benchmark_format=json")
    assert r.returncode == 0, r.stderr
    expect = ["BMMatrixDouble/1", "BMMatrixDouble/8", "BMMatrixDouble/30", "BMMatrixSingle/1", "BMMatrixSingle/8", "BMMatrixSingle/30"]
    check_schema(json.loads(r.stdout), expect)
    check_schema(json.load(open(out)), expect)


def test_loads_a_real_file_when_present(tmp_path):
    os.makedirs(tmp_path / "moderate")
    host.write_matrix_market(tmp_path / "moderate" / "ecology2.mtx", host.gen_laplace2d(30))
    r = run("--cpu", "--iterations=1", "--warmup=0", "--benchmark_filter=Double/8$", "--mat-dir=%s" % tmp_path, "--benchmark_format=json")
    b = json.loads(r.stdout)["benchmarks"][0]
    assert b["rows"] == 900 and b["stored"] == 5 * 900 - 4 * 30 and b["source"].endswith("ecology2.mtx")


def test_banded_generator_shape():
    a = host.gen_banded(5000, 17, 300)
    rp, ci = a.rowptr, a.colidx
    assert a.m == a.n == 5000 and np.all(np.diff(rp) == 17)
    rows = np.repeat(np.arange(a.m), np.diff(rp))
    assert np.all(np.abs(ci - rows) <= 300)
    for i in (0, 1, 2499, 4999):
        c = ci[rp[i]:rp[i + 1]]
        assert np.all(np.diff(c) > 0) and i in c
    diag = a.values[ci == rows]
    off = np.abs(np.where(ci == rows, 0.0, a.values))
    assert np.all(diag > np.add.reduceat(off, rp[:-1]))  # strictly diagonally dominant


@pytest.mark.gpu
def test_gpu_path_runs_and_reports(tmp_path):
    out = tmp_path / "o.json"
    r = run("--scale=0.05", "--iterations=5", "--benchmark_filter=/(0|1|8|23)$", "--benchmark_out=%s" % out)
    assert r.returncode == 0, r.stderr
    doc = json.load(open(out))
    check_schema(doc, ["BMMatrixDouble/0", "BMMatrixDouble/1", "BMMatrixDouble/8", "BMMatrixDouble/23",
                       "BMMatrixSingle/0", "BMMatrixSingle/1", "BMMatrixSingle/8", "BMMatrixSingle/23"])
    assert doc["context"]["path"] == "gpu" and all(b["real_time"] < 50 for b in doc["benchmarks"])
    assert "BMMatrixDouble/23" in r.stdout and "big/atmosmodd" in r.stdout
