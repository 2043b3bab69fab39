"""The C++ harnesses (respasol_b200/bin/test_spmv, test_ilu0) keep the reference's command line and stdout
(GPU/spmv.cu:35-42,203-207,260; GPU/ilu0.cu:221-226,278-282,313-317) -- run them as a user would."""
import os
import re
import subprocess

import pytest

from respasol_b200 import host

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "respasol_b200", "bin")
GOLD = os.path.join(ROOT, "tests", "golden")


def run(exe, *args):
    r = subprocess.run([os.path.join(BIN, exe), *map(str, args)], capture_output=True, text=True, timeout=300)
    return r.returncode, r.stdout, r.stderr


@pytest.fixture(scope="module")
def lap_mtx(tmp_path_factory):
    p = tmp_path_factory.mktemp("mtx") / "lap200.mtx"
    host.write_matrix_market(p, host.gen_laplace2d(200))
    return p


def test_usage_without_arguments():
    for exe in ("test_spmv", "test_ilu0"):
        rc, out, err = run(exe)
        assert rc != 0 and "sage" in (out + err)  # "Usage: ..." and return -1, like spmv.cu:35-42


@pytest.mark.parametrize("flag,word", [("--precision=d", "DOUBLE"), ("--precision=s", "SINGLE"), ("--precision=m", "DOUBLE")])
def test_spmv_stdout_contract(lap_mtx, flag, word):
    rc, out, err = run("test_spmv", lap_mtx, flag)
    assert rc == 0, err
    lines = out.strip().splitlines()
    m = re.fullmatch(word + r" PRECISION SPMV solve time \(microseconds\) = (\d+\.\d{6})", lines[0])
    assert m and 0.5 < float(m.group(1)) < 5000
    m = re.fullmatch(r"Error= (\d\.\d{6}e[+-]\d{2})", lines[1])
    assert m and float(m.group(1)) <= (1e-12 if "=s" not in flag else 1e-5)


def test_spmv_on_reference_fixture():
    rc, out, err = run("test_spmv", os.path.join(GOLD, "b1_ss.mtx"))
    assert rc == 0 and out.startswith("DOUBLE PRECISION SPMV solve time (microseconds) = ") and "\nError= " in out


@pytest.mark.parametrize("flag,word", [("--precision=d", "DOUBLE"), ("--precision=s", "SINGLE")])
def test_ilu0_stdout_contract(lap_mtx, flag, word):
    rc, out, err = run("test_ilu0", lap_mtx, flag)
    assert rc == 0, err
    pat = (word + r" PRECISION SOLVE IN  MILLISECONDS\n Symbolic = (\d+\.\d{6})\n Numeric = (\d+\.\d{6}) \n"
           r" Symbolic\+ Numeric = (\d+\.\d{6})\n Solve = (\d+\.\d{6})\n")
    m = re.search(pat, out)
    assert m, out
    sym, num, both, solve = map(float, m.groups())
    assert abs(both - (sym + num)) < 1e-3 and 0 < num < 1000 and 0 < solve < 1000


def test_ilu0_zero_pivot_messages(tmp_path):
    rc, out, err = run("test_ilu0", os.path.join(GOLD, "b1_ss.mtx"))
    assert rc == 0 and out.startswith("A(0,0) is missing\n")  # ilu0.cu:221-226: print and return 0
    p = tmp_path / "sing.mtx"
    p.write_text("%%MatrixMarket matrix coordinate real general\n2 2 4\n1 1 1\n1 2 1\n2 1 1\n2 2 1\n")
    rc, out, err = run("test_ilu0", p)
    assert rc == 0 and "L(1,1) is zero\n" in out              # ilu0.cu:278-282
