"""Our Matrix-Market loader (respasol_b200/host/rsg_loader.cpp) against the reference's own loader:
(1) committed golden vectors produced by tests/golden/make_golden.py from the reference code,
(2) live against oracle/_ref where it was built, on generated files."""
import json
import os

import numpy as np
import pytest

import orc
from respasol_b200 import host

GOLD = os.path.join(os.path.dirname(__file__), "golden")
with open(os.path.join(GOLD, "loader_golden.json")) as f:
    GOLDEN = json.load(f)


@pytest.mark.parametrize("key", sorted(GOLDEN))
def test_loader_matches_reference_golden(key):
    name, b, t = key.split("|")
    base, tr = int(b.split("=")[1]), int(t.split("=")[1])
    g = GOLDEN[key]
    path = os.path.join(GOLD, name)
    if g["ok"] != 1:
        with pytest.raises(ValueError):
            host.load_matrix_market(path, base, tr)
        return
    a = host.load_matrix_market(path, base, tr)
    assert (a.m, a.n, a.nnz, a.is_symmetric) == (g["m"], g["n"], g["nnz"], g["isSymmetric"])
    assert a.rowptr.tolist() == g["rowptr"]
    assert a.colidx.tolist() == g["colidx"]
    assert [float(v).hex() for v in a.values] == g["values_hex"]


def test_b1ss_kat():
    a = host.load_matrix_market(os.path.join(GOLD, "b1_ss.mtx"), 1, 0)
    assert a.rowptr.tolist() == [1, 4, 6, 8, 10, 12, 14, 16]
    t = host.load_matrix_market(os.path.join(GOLD, "b1_ss.mtx"), 0, 1)
    assert t.colidx[:2].tolist() == [4, 5] and t.values[0] == -.03599942


def test_symmetric_fixed_mode_expands():
    a = host.load_matrix_market(os.path.join(GOLD, "bcspwr01.mtx"), 0, 0, host.LOADER_EXPAND_SYMMETRIC)
    assert a.nnz == 131 and a.stored == 131
    # symmetric pattern after expansion
    d = np.zeros((39, 39))
    for i in range(39):
        d[i, a.colidx[a.rowptr[i]:a.rowptr[i + 1]]] = 1
    assert np.array_equal(d, d.T)


def test_roundtrip_writer(tmp_path):
    a = host.gen_laplace2d(12)
    p = tmp_path / "lap.mtx"
    host.write_matrix_market(p, a)
    b = host.load_matrix_market(p, 0, 0)
    assert np.array_equal(a.rowptr, b.rowptr) and np.array_equal(a.colidx, b.colidx) and np.array_equal(a.values, b.values)


def test_missing_file_does_not_exit(tmp_path):
    with pytest.raises(ValueError):
        host.load_matrix_market(tmp_path / "nope.mtx")


@pytest.mark.skipif(not orc.have_ref(), reason="oracle/_ref not built (reference tree absent)")
@pytest.mark.parametrize("base,tr", [(0, 0), (1, 0), (0, 1)])
def test_live_against_reference_loader(tmp_path, base, tr):
    a = host.gen_femlike(6, 0.6, 11)
    a.values = a.values * (1 + 1e-7)  # non-trivial digits through %.17g
    p = tmp_path / "fem.mtx"
    host.write_matrix_market(p, a)
    r = orc.ref_load(p, base, tr)
    b = host.load_matrix_market(p, base, tr)
    assert r["ok"] == 1 and (b.m, b.n, b.nnz) == (r["m"], r["n"], r["nnz"])
    assert np.array_equal(b.rowptr, r["rowptr"]) and np.array_equal(b.colidx, r["colidx"])
    assert np.array_equal(b.values, r["values"])
