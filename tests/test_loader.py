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


# ---- parallel ingest: files above 1 MB are tokenised by all OpenMP threads; the result must be the
# sequential tokeniser's -- and the reference loader's -- bit for bit, and anything irregular must fall back.
def _write_big(path, banner, entries_lines, m, n, nnz):
    with open(path, "w") as f:
        f.write(banner + "\n%% a comment line\n%d %d %d\n" % (m, n, nnz))
        f.write("".join(entries_lines))


def _big_entries(N=330, seed=5, fmt="%d %d %.17g\n", shuffle=True, dup=True):
    a = host.gen_laplace2d(N)
    rng = np.random.default_rng(seed)
    rows = np.repeat(np.arange(a.m), np.diff(a.rowptr)) + 1
    cols = a.colidx + 1
    vals = a.values * (1 + 1e-9 * rng.standard_normal(a.stored))
    if dup:  # duplicate coordinates exercise the reference's unstable per-row sort
        k = rng.integers(0, a.stored, 2000)
        rows, cols, vals = np.concatenate([rows, rows[k]]), np.concatenate([cols, cols[k]]), np.concatenate([vals, -vals[k]])
    if shuffle:
        p = rng.permutation(len(rows))
        rows, cols, vals = rows[p], cols[p], vals[p]
    return a.m, a.n, rows, cols, vals


def _same(a, b):
    return (a.m, a.n, a.nnz) == (b.m, b.n, b.nnz) and np.array_equal(a.rowptr, b.rowptr) and \
        np.array_equal(a.colidx[:a.stored], b.colidx[:b.stored]) and \
        np.array_equal(a.values[:a.stored].view(np.uint64), b.values[:b.stored].view(np.uint64))


def _load_both(path, base=0, tr=0):
    par = host.load_matrix_market(path, base, tr)
    os.environ["RSG_LOADER_THREADS"] = "1"
    try:
        seq = host.load_matrix_market(path, base, tr)
    finally:
        del os.environ["RSG_LOADER_THREADS"]
    return par, seq


@pytest.mark.parametrize("variant", ["regular", "crlf_blank_lines", "split_entry", "zero_based", "pattern", "integer", "tabs"])
@pytest.mark.parametrize("base,tr", [(0, 0), (1, 1)])
def test_parallel_ingest_equals_sequential_and_reference(tmp_path, variant, base, tr):
    m, n, rows, cols, vals = _big_entries()
    banner = "%%MatrixMarket matrix coordinate real general"
    if variant == "pattern":
        banner = "%%MatrixMarket matrix coordinate pattern general"
        lines = ["%d %d\n" % (r, c) for r, c in zip(rows, cols)]
    elif variant == "integer":
        banner = "%%MatrixMarket matrix coordinate integer general"
        lines = ["%d %d %d\n" % (r, c, int(v * 1000)) for r, c, v in zip(rows, cols, vals)]
    else:
        lines = ["%d %d %.17g\n" % (r, c, v) for r, c, v in zip(rows, cols, vals)]
    if variant == "crlf_blank_lines":
        lines = [ln[:-1] + "\r\n" + ("\n" if i % 1000 == 0 else "") for i, ln in enumerate(lines)]
    if variant == "tabs":
        lines = ["  " + ln.replace(" ", "\t") for ln in lines]
    if variant == "split_entry":  # one entry broken across two lines: legal for fscanf, so the parallel pass must give up
        k = len(lines) // 2
        r, c, v = lines[k].split()
        lines[k] = "%s %s\n%s\n" % (r, c, v)
    if variant == "zero_based":
        lines = ["%d %d %.17g\n" % (r - 1, c - 1, v) for r, c, v in zip(rows, cols, vals)]
    p = tmp_path / "big.mtx"
    _write_big(p, banner, lines, m, n, len(rows))
    assert os.path.getsize(p) > (1 << 20)
    par, seq = _load_both(p, base, tr)
    assert _same(par, seq)
    if orc.have_ref():
        r = orc.ref_load(p, base, tr)
        assert r["ok"] == 1 and np.array_equal(par.rowptr, r["rowptr"]) and np.array_equal(par.colidx[:par.stored], r["colidx"])
        assert np.array_equal(par.values[:par.stored], r["values"])


@pytest.mark.parametrize("defect", ["count_too_small", "count_too_large", "out_of_range", "garbage_token"])
def test_parallel_ingest_error_paths_match_sequential(tmp_path, defect):
    m, n, rows, cols, vals = _big_entries(dup=False)
    lines = ["%d %d %.17g\n" % (r, c, v) for r, c, v in zip(rows, cols, vals)]
    nnz = len(lines)
    if defect == "count_too_small":
        nnz -= 3
    elif defect == "count_too_large":
        nnz += 3
    elif defect == "out_of_range":
        lines[len(lines) // 3] = "%d %d 1.0\n" % (m + 1, 1)
    else:
        lines[len(lines) // 3] = "12 x 1.0\n"
    p = tmp_path / "bad.mtx"
    _write_big(p, "%%MatrixMarket matrix coordinate real general", lines, m, n, nnz)
    for threads in (None, "1"):
        if threads:
            os.environ["RSG_LOADER_THREADS"] = threads
        try:
            with pytest.raises(ValueError):
                host.load_matrix_market(p)
        finally:
            os.environ.pop("RSG_LOADER_THREADS", None)
