"""SpMV parity: the CUDA path through the C ABI against the CPU oracle on the same seeded inputs.

Tolerances (stated by BASELINE.json north_star / SURVEY.md 8c):
  fp64:   ||y - y_ref||inf <= 1e-12 * ||y_ref||inf
  fp32 / mixed: per row |dy_i| <= (len_i + 2) * eps * sum_k |a_ik x_k|, eps = 2^-24 (fp32) or
                2^-53 (mixed accumulates in fp64) -- the ulp-scaled bound for any summation order.
"""
import os

import numpy as np
import pytest
import torch

import orc
from respasol_b200 import host
from respasol_b200.gpu import ALGO, DeviceCSR, Handle

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")


def row_abs_sums(a, x, precision):
    vt = np.float64 if precision == "d" else np.float32
    xt = np.float32 if precision == "s" else np.float64
    absa = host.CSR(a.m, a.n, a.nnz, a.rowptr, a.colidx, np.abs(a.values.astype(vt)).astype(np.float64), a.base)
    return orc.spmv(absa, np.abs(x.astype(xt)).astype(np.float64), "d")


def check(handle, a, x, precision, alpha=1.0, beta=0.0, y0=None):
    A = DeviceCSR(a, precision)
    xt = torch.float32 if precision == "s" else torch.float64
    xd = torch.from_numpy(x).to("cuda").to(xt)
    yd = None if y0 is None else torch.from_numpy(y0).to("cuda").to(xt)
    y = handle.spmv(A, xd, yd, alpha, beta).cpu().numpy()
    ref = orc.spmv(a, x, precision, alpha, beta, y0)
    if precision == "d":
        scale = max(np.abs(ref).max(), 1e-300) if a.m else 1.0
        assert np.abs(y - ref).max(initial=0.0) <= 1e-12 * scale
    else:
        eps = 2.0 ** -24 if precision == "s" else 2.0 ** -53
        lens = np.diff(a.rowptr).astype(np.float64)
        bound = (lens + 2) * eps * (abs(alpha) * row_abs_sums(a, x, precision) + abs(beta) * (0 if y0 is None else np.abs(y0)))
        if precision == "m":  # output rounding of the fp64 result is below the bound already
            bound = bound + 1e-300
        assert np.all(np.abs(y.astype(np.float64) - ref.astype(np.float64)) <= bound + 1e-300)
    return y


SMALL = {
    "one": lambda: host.load_matrix_market(os.path.join(GOLD, "one.mtx")),
    "b1_ss": lambda: host.load_matrix_market(os.path.join(GOLD, "b1_ss.mtx")),
    "b1_ss_base1": lambda: host.load_matrix_market(os.path.join(GOLD, "b1_ss.mtx"), 1),
    "bcspwr01": lambda: host.load_matrix_market(os.path.join(GOLD, "bcspwr01.mtx")),
    "rect": lambda: host.load_matrix_market(os.path.join(GOLD, "rect.mtx")),
    "empty_rows": lambda: host.load_matrix_market(os.path.join(GOLD, "empty_rows.mtx")),
    "dup_unsorted": lambda: host.load_matrix_market(os.path.join(GOLD, "dup_unsorted.mtx")),
    "lap2d_64": lambda: host.gen_laplace2d(64),
    "stencil27_20": lambda: host.gen_stencil27(20, 20, 20, seed=2),
    "stencil27_slab": lambda: host.gen_stencil27(12, 12, 30, 10, 20, seed=2),
    "femlike_12": lambda: host.gen_femlike(12),
    "powerlaw_20k": lambda: host.gen_powerlaw(20000, 2.2, 64, 2, 9000, 4),
}


@pytest.mark.parametrize("name", sorted(SMALL))
@pytest.mark.parametrize("precision", ["d", "s", "m"])
@pytest.mark.parametrize("algo", ["auto", "window", "tile", "rowblock", "merge", "vector", "scalar"])
def test_spmv_small(handle, name, precision, algo):
    a = SMALL[name]()
    handle.set_option("SPMV_ALGO", ALGO[algo])
    try:
        x = orc.dlarnv(a.n) - 0.5
        check(handle, a, x, precision)
    finally:
        handle.set_option("SPMV_ALGO", 0)


def test_golden_b1ss_values(handle):
    a = SMALL["b1_ss"]()
    y = check(handle, a, np.ones(7), "d")
    np.testing.assert_allclose(y, [3, -0.55, -0.9, -0.55, 0.96400058, 0.9823629, 0.992278221], rtol=1e-15)
    y = check(handle, a, orc.dlarnv(7), "d")
    np.testing.assert_allclose(y, [1.1964668762652444, -0.50611051611001145, 0.019299642814089787, -0.041548033327566514,
                                   0.30173623575092434, 0.81428611598229961, 0.99624904350899124], rtol=4e-16)


@pytest.mark.parametrize("precision", ["d", "s", "m"])
@pytest.mark.parametrize("algo", ["window", "tile", "rowblock", "merge", "vector"])
def test_alpha_beta(handle, precision, algo):
    a = host.gen_stencil27(9, 10, 11, seed=5)
    handle.set_option("SPMV_ALGO", ALGO[algo])
    try:
        x = orc.dlarnv(a.n) - 0.5
        y0 = (orc.dlarnv(a.m + 3)[3:] - 0.5)
        check(handle, a, x, precision, alpha=-1.75, beta=0.5, y0=y0)
        check(handle, a, x, precision, alpha=0.0, beta=2.0, y0=y0)
    finally:
        handle.set_option("SPMV_ALGO", 0)


def test_beta_zero_ignores_nan_y(handle):
    a = host.gen_laplace2d(30)
    A = DeviceCSR(a, "d")
    x = torch.ones(a.n, dtype=torch.float64, device="cuda")
    y = torch.full((a.m,), float("nan"), dtype=torch.float64, device="cuda")
    handle.spmv(A, x, y)
    assert torch.isfinite(y).all()


@pytest.mark.parametrize("variant", [0, 1, 2, 3, 4])
@pytest.mark.parametrize("reduce", [0, 1, 2, 4, 8, 16, 32])
def test_rowblock_variants(handle, variant, reduce):
    a = host.gen_femlike(14, 0.7, 9)
    handle.set_option("SPMV_ALGO", ALGO["rowblock"])
    handle.set_option("SPMV_VARIANT", variant)
    handle.set_option("SPMV_REDUCE", reduce)
    try:
        check(handle, a, orc.dlarnv(a.n) - 0.5, "d")
        check(handle, a, orc.dlarnv(a.n) - 0.5, "s")
    finally:
        handle.set_option("SPMV_ALGO", 0)
        handle.set_option("SPMV_VARIANT", 0)
        handle.set_option("SPMV_REDUCE", 0)


WINDOWED = {
    "lap2d_37": lambda: host.gen_laplace2d(37),
    "stencil27_11": lambda: host.gen_stencil27(11, 12, 13, seed=4),
    "stencil27_slab": lambda: host.gen_stencil27(12, 12, 30, 10, 20, seed=2),
    "femlike_14": lambda: host.gen_femlike(14, 0.7, 9),
    "b1_ss_base1": lambda: host.load_matrix_market(os.path.join(GOLD, "b1_ss.mtx"), 1),
    "bcspwr01": lambda: host.load_matrix_market(os.path.join(GOLD, "bcspwr01.mtx")),
    "rect": lambda: host.load_matrix_market(os.path.join(GOLD, "rect.mtx")),
    "empty_rows": lambda: host.load_matrix_market(os.path.join(GOLD, "empty_rows.mtx")),
}


@pytest.mark.parametrize("variant", [0, 1, 2])
@pytest.mark.parametrize("reduce", [0, 1, 2, 4, 8, 16, 32])
@pytest.mark.parametrize("win_bytes", [0, 8192, 20000, 90000])
def test_window_variants(handle, variant, reduce, win_bytes):
    """every CTA width x every lanes-per-row reduction x tile budgets (many small tiles ... few fat ones) of the
    x-window kernel; thread-per-row sums left to right like the oracle: bit-exact"""
    handle.set_option("SPMV_ALGO", ALGO["window"])
    handle.set_option("SPMV_VARIANT", variant)
    handle.set_option("SPMV_REDUCE", reduce)
    handle.set_option("SPMV_WIN_BYTES", win_bytes)
    try:
        for name in sorted(WINDOWED):
            a = WINDOWED[name]()
            x = orc.dlarnv(a.n) - 0.5
            for p in ("d", "s", "m"):
                check(handle, a, x, p)
            if a.m >= 64:
                assert handle.info("SPMV_ALGO_USED") == ALGO["window"], name
            if reduce == 1:
                A = DeviceCSR(a, "d")
                y = handle.spmv(A, torch.from_numpy(x).cuda()).cpu().numpy()
                assert np.array_equal(y, orc.spmv(a, x, "d")), name
    finally:
        handle.set_option("SPMV_ALGO", 0)
        handle.set_option("SPMV_VARIANT", 0)
        handle.set_option("SPMV_REDUCE", 0)
        handle.set_option("SPMV_WIN_BYTES", 0)


def test_window_ragged_ends_and_views(handle):
    """x lengths that are not multiples of 4 (the last entries of x travel by plain loads), entry counts that are not
    multiples of 8 (the last entries of a tile too), row-slab views whose first entry is not entry 0, a gap setting
    that merges every run into one, and an x view that is not 16-byte aligned (falls back to the tile plan)."""
    handle.set_option("SPMV_ALGO", ALGO["window"])
    try:
        for n in (61, 62, 63, 64, 65, 67, 130):
            rng = np.random.default_rng(n)
            m = n + 3
            rows = []
            for i in range(m):  # banded rows that reach the last columns, ragged lengths
                c0 = min(max(0, i - 3), n - 1)
                cs = sorted(set(range(c0, min(n, c0 + 1 + int(rng.integers(0, 7))))) | {n - 1 - (i % 3)})
                rows.append(cs)
            rp = np.zeros(m + 1, dtype=np.int32); rp[1:] = np.cumsum([len(r) for r in rows])
            ci = np.fromiter((c for r in rows for c in r), dtype=np.int32, count=int(rp[-1]))
            a = host.CSR(m, n, int(rp[-1]), rp, ci, rng.uniform(-1, 1, int(rp[-1])), 0)
            x = orc.dlarnv(n) - 0.5
            for p in ("d", "s", "m"):
                check(handle, a, x, p, alpha=1.25, beta=-0.5, y0=orc.dlarnv(m) - 0.5)
            assert handle.info("SPMV_ALGO_USED") == ALGO["window"]
        a = host.gen_stencil27(12, 12, 30, 10, 20, seed=2)
        x = orc.dlarnv(a.n) - 0.5
        A = DeviceCSR(a, "d")
        xd = torch.from_numpy(x).cuda()
        ref = orc.spmv(a, x, "d")
        for r0, r1 in ((0, a.m), (5, 700), (333, a.m), (1001, 1002)):
            y = handle.spmv(A.view_rows(r0, r1), xd).cpu().numpy()
            assert np.array_equal(y, ref[r0:r1]), (r0, r1)
        for gap in (1, 100000):
            handle.set_option("SPMV_WIN_GAP", gap)
            assert np.array_equal(handle.spmv(DeviceCSR(a, "d"), xd).cpu().numpy(), ref)
        handle.set_option("SPMV_WIN_GAP", 0)
        xo = torch.empty(a.n + 1, dtype=torch.float64, device="cuda")
        xo[1:] = xd
        y = handle.spmv(A, xo[1:]).cpu().numpy()
        assert np.abs(y - ref).max() <= 1e-12 * np.abs(ref).max()
        assert handle.info("SPMV_ALGO_USED") == ALGO["tile"]
    finally:
        handle.set_option("SPMV_ALGO", 0)
        handle.set_option("SPMV_WIN_GAP", 0)


def test_window_declines_scattered_patterns(handle):
    """hashed columns / dense rows have no compact x windows: the plan falls back (AUTO: vector or tile), results unchanged"""
    a = host.gen_powerlaw(20000, 2.2, 64, 2, 9000, 4)
    x = orc.dlarnv(a.n) - 0.5
    for algo in ("window", "auto"):
        handle.set_option("SPMV_ALGO", ALGO[algo])
        check(handle, a, x, "d")
        assert handle.info("SPMV_ALGO_USED") != ALGO["window"]
    handle.set_option("SPMV_ALGO", 0)


@pytest.mark.parametrize("variant", list(range(10)))
@pytest.mark.parametrize("reduce", [0, 1, 4, 32])
@pytest.mark.parametrize("rpt", [0, 3])
def test_tile_variants(handle, variant, reduce, rpt):
    """every compiled shape of the tile kernel (one tile per CTA): fitting tiles, multi-pass tiles (rows longer than
    the tile slack), CTA-per-row rows, a ragged last quad.  (The retired pipe / stream enum values select it too.)"""
    handle.set_option("SPMV_ALGO", ALGO["pipe" if variant == 0 and rpt == 3 else "stream" if variant == 0 else "tile"])
    handle.set_option("SPMV_VARIANT", variant)
    handle.set_option("SPMV_REDUCE", reduce)
    handle.set_option("SPMV_ROWS_PER_THREAD", rpt)
    try:
        mats = [host.gen_femlike(14, 0.7, 9), host.gen_powerlaw(16000, 2.2, 48, 3, min(700 + 1733 * variant, 15000), 4),
                host.gen_laplace2d(37)]
        for a in mats:
            x = orc.dlarnv(a.n) - 0.5
            for p in ("d", "s", "m"):
                check(handle, a, x, p)
            assert handle.info("SPMV_ALGO_USED") == ALGO["tile"]
        if reduce == 1:  # thread-per-row sums left to right like the oracle: bit-exact
            a = host.gen_stencil27(11, 12, 13, seed=4)
            x = orc.dlarnv(a.n) - 0.5
            A = DeviceCSR(a, "d")
            y = handle.spmv(A, torch.from_numpy(x).cuda()).cpu().numpy()
            assert np.array_equal(y, orc.spmv(a, x, "d"))
    finally:
        handle.set_option("SPMV_ALGO", 0)
        handle.set_option("SPMV_VARIANT", 0)
        handle.set_option("SPMV_REDUCE", 0)
        handle.set_option("SPMV_ROWS_PER_THREAD", 0)


def test_pipe_tiny_tiles_and_slabs(handle):
    """tile size forced down to 64 items: many tiles per CTA, empty tiles inside long rows, slab views
    whose first entry is not entry 0"""
    _tiny_tiles(handle, "tile")


def _tiny_tiles(handle, algo):
    handle.set_option("SPMV_ALGO", ALGO[algo])
    handle.set_option("SPMV_LONG_ROW", 64)
    try:
        for a in (host.gen_stencil27(12, 12, 30, 10, 20, seed=2), host.gen_powerlaw(3000, 2.2, 48, 2, 2500, 4),
                  SMALL["empty_rows"](), SMALL["rect"](), SMALL["b1_ss_base1"]()):
            x = orc.dlarnv(a.n) - 0.5
            for p in ("d", "s", "m"):
                check(handle, a, x, p, alpha=0.75, beta=-0.5, y0=orc.dlarnv(a.m) - 0.25)
    finally:
        handle.set_option("SPMV_ALGO", 0)
        handle.set_option("SPMV_LONG_ROW", 0)


@pytest.mark.parametrize("hint", [0, 1, 2])
@pytest.mark.parametrize("persist", [0, 1])
def test_cache_hints(handle, hint, persist):
    a = host.gen_stencil27(16, 16, 16, seed=2)
    handle.set_option("SPMV_CACHE_HINT", hint)
    handle.set_option("SPMV_X_PERSIST", persist)
    try:
        for algo in ("window", "tile", "rowblock", "merge"):
            handle.set_option("SPMV_ALGO", ALGO[algo])
            check(handle, a, orc.dlarnv(a.n), "d")
            check(handle, a, orc.dlarnv(a.n), "m")
    finally:
        handle.set_option("SPMV_CACHE_HINT", 1)
        handle.set_option("SPMV_X_PERSIST", 0)
        handle.set_option("SPMV_X_HINT", 0)
        handle.set_option("SPMV_ALGO", 0)


def test_long_rows_cta_per_row_and_merge(handle):
    # rows far longer than a tile: the row-block kernel falls to CTA-per-row, AUTO picks the tile kernel with split rows
    a = host.gen_powerlaw(60000, 2.2, 32, 3, 50000, 7)
    x = orc.dlarnv(a.n) - 0.5
    for algo in ("window", "rowblock", "merge", "tile", "auto"):
        handle.set_option("SPMV_ALGO", ALGO[algo])
        for p in ("d", "s", "m"):
            check(handle, a, x, p)
    handle.set_option("SPMV_ALGO", 0)
    check(handle, a, x, "d")
    assert handle.info("SPMV_ALGO_USED") == ALGO["tile"]
    assert handle.info("SPMV_MAX_ROW") == int(np.diff(a.rowptr).max())


def test_unaligned_views_take_scalar_path(handle):
    a = host.gen_stencil27(10, 10, 10, seed=3)
    A = DeviceCSR(a, "d")
    # shift colidx/values by one element so the pointers are not 16-byte aligned
    ci = torch.empty(A.colidx.numel() + 1, dtype=torch.int32, device="cuda")
    va = torch.empty(A.values.numel() + 1, dtype=torch.float64, device="cuda")
    ci[1:] = A.colidx; va[1:] = A.values
    A.colidx, A.values = ci[1:], va[1:]
    x = orc.dlarnv(a.n)
    ref = orc.spmv(a, x)
    for algo in ("auto", "window", "tile", "rowblock", "merge"):  # window / tile fall back to rowblock's scalar staging
        handle.set_option("SPMV_ALGO", ALGO[algo])
        y = handle.spmv(A, torch.from_numpy(x).cuda()).cpu().numpy()
        assert np.abs(y - ref).max() <= 1e-12 * np.abs(ref).max()
    handle.set_option("SPMV_ALGO", 0)


def test_empty_and_degenerate(handle):
    z = host.CSR(0, 0, 0, np.zeros(1, np.int32), np.zeros(0, np.int32), np.zeros(0))
    Z = DeviceCSR(z, "d")
    Z.colidx = torch.zeros(1, dtype=torch.int32, device="cuda"); Z.values = torch.zeros(1, dtype=torch.float64, device="cuda")
    handle.spmv(Z, torch.zeros(1, dtype=torch.float64, device="cuda"), torch.zeros(1, dtype=torch.float64, device="cuda"))
    # all rows empty
    e = host.CSR(5, 5, 0, np.zeros(6, np.int32), np.zeros(0, np.int32), np.zeros(0))
    E = DeviceCSR(e, "d")
    E.colidx = torch.zeros(1, dtype=torch.int32, device="cuda"); E.values = torch.zeros(1, dtype=torch.float64, device="cuda")
    for algo in ("window", "tile", "rowblock", "merge", "vector", "scalar"):
        handle.set_option("SPMV_ALGO", ALGO[algo])
        y = handle.spmv(E, torch.ones(5, dtype=torch.float64, device="cuda"),
                        torch.full((5,), 7.0, dtype=torch.float64, device="cuda"))
        assert y.cpu().tolist() == [0.0] * 5
    handle.set_option("SPMV_ALGO", 0)


def test_host_pointer_entry_points(handle):
    a = host.gen_femlike(10)
    x = orc.dlarnv(a.n) - 0.5
    for p in ("d", "s", "m"):
        y = handle.spmv_host(a, x, p)
        ref = orc.spmv(a, x, p)
        np.testing.assert_allclose(y, ref, rtol=1e-12 if p != "s" else 2e-5, atol=1e-12 if p != "s" else 1e-5)
    for p in ("d", "s", "m"):
        mat = handle.matrix_upload(a, p)
        xt = np.float32 if p == "s" else np.float64
        y = np.zeros(a.m, dtype=xt)
        handle.matrix_spmv_host(mat, x.astype(xt), y)
        ref = orc.spmv(a, x, p)
        np.testing.assert_allclose(y, ref, rtol=1e-12 if p != "s" else 2e-5, atol=1e-12 if p != "s" else 1e-5)
        handle.matrix_free(mat)


def long_rows_matrix(n, lens, rows_at, base=0, seed=11):
    """diagonal + a few long rows of the given lengths (contiguous column runs starting at varying offsets)"""
    rng = np.random.default_rng(seed)
    cols = [[i] for i in range(n)]
    for r, ln in zip(rows_at, lens):
        start = int(rng.integers(0, n - ln + 1))
        cols[r] = sorted(set(range(start, start + ln)) | {r})
    rp = np.zeros(n + 1, dtype=np.int32)
    rp[1:] = np.cumsum([len(c) for c in cols])
    ci = np.fromiter((c for row in cols for c in row), dtype=np.int32, count=int(rp[-1]))
    v = rng.uniform(-1, 1, int(rp[-1]))
    return host.CSR(n, n, int(rp[-1]), rp + base, ci + base, v, base)


@pytest.mark.parametrize("precision", ["d", "s", "m"])
@pytest.mark.parametrize("algo", ["auto", "tile"])
def test_long_rows_are_split_into_segments(handle, precision, algo):
    """Rows longer than a CTA's stage (2304 / 4352 entries) are cut into 2048-entry segments, one CTA each, whose
    partial sums the last arrival adds in order: lengths around every boundary, first/last row, neighbours."""
    n = 30000
    lens = [2304, 2305, 2309, 4352, 4353, 4357, 6143, 6144, 6145, 8193, 20000, 29999, 4097, 2049 + 2304]
    rows_at = [0, 1, 2, 77, 78, 500, 501, 9000, 9001, 15000, 15001, 29998, 29999, 12345]
    handle.set_option("SPMV_ALGO", ALGO[algo])
    try:
        for base in (0, 1):
            a = long_rows_matrix(n, lens, rows_at, base)
            x = orc.dlarnv(a.n) - 0.5
            check(handle, a, x, precision)
            assert handle.info("SPMV_ALGO_USED") == ALGO["tile"]
            y0 = orc.dlarnv(a.m) - 0.5
            check(handle, a, x, precision, alpha=-1.5, beta=0.25, y0=y0)
            check(handle, a, x, precision)  # counters were reset by the previous products
    finally:
        handle.set_option("SPMV_ALGO", 0)


def test_long_row_split_is_deterministic(handle):
    a = host.gen_powerlaw(60000, 2.2, 64, 3, 20000, 4)
    A = DeviceCSR(a, "d")
    x = torch.from_numpy(orc.dlarnv(a.n) - 0.5).cuda()
    y0 = handle.spmv(A, x).clone()
    assert handle.info("SPMV_ALGO_USED") == ALGO["tile"]
    for _ in range(5):
        assert torch.equal(handle.spmv(A, x), y0)


def test_too_many_long_rows_fall_back_to_merge_path(handle):
    n, ln = 4300, 4400  # more than 4096 rows longer than any stage
    rp = (np.arange(n + 1, dtype=np.int64) * ln).astype(np.int32)
    ci = ((np.arange(n)[:, None] * 7 + np.arange(ln)[None, :]) % n).astype(np.int32)
    ci.sort(axis=1)
    rng = np.random.default_rng(3)
    a = host.CSR(n, n, n * ln, rp, ci.ravel().copy(), rng.uniform(-1, 1, n * ln), 0)
    x = orc.dlarnv(a.n) - 0.5
    check(handle, a, x, "d")
    assert handle.info("SPMV_ALGO_USED") == ALGO["merge"]
    handle.set_option("SPMV_ALGO", ALGO["tile"])  # forced: long rows stay CTA-per-row inside the tiles
    try:
        check(handle, a, x, "d")
    finally:
        handle.set_option("SPMV_ALGO", 0)


def test_auto_picks_the_unstaged_kernel_for_scattered_columns(handle):
    """Rows whose neighbours gather unrelated x lines (hashed columns inside a band) take the plain sub-warp-per-row
    kernel (all of L1 for x), and so do small matrices without dense rows (the tile kernel's per-CTA set-up shows in a
    10 us product); large matrices with neighbour coherence and anything with dense rows stay tiled."""
    x_of = lambda a: orc.dlarnv(a.n) - 0.5
    a = host.gen_banded(60000, 40, 6000)
    for p in ("d", "s", "m"):
        check(handle, a, x_of(a), p)
        assert handle.info("SPMV_ALGO_USED") == ALGO["vector"]
    check(handle, a, x_of(a), "d", alpha=-0.5, beta=2.0, y0=orc.dlarnv(a.m))
    b = host.gen_banded(60000, 8, 6000)
    check(handle, b, x_of(b), "d")
    assert handle.info("SPMV_ALGO_USED") == ALGO["vector"]
    for b in (host.gen_stencil27(30, seed=2), host.gen_femlike(24), host.gen_stencil27(60, seed=2)):  # neighbour coherence: x windows
        check(handle, b, x_of(b), "d")
        assert handle.info("SPMV_ALGO_USED") == ALGO["window"]
    b = host.gen_powerlaw(60000, 2.2, 64, 3, 20000, 4)
    check(handle, b, x_of(b), "d")
    assert handle.info("SPMV_ALGO_USED") == ALGO["tile"]
    # one row far longer than the rest in a small matrix: the sub-warp kernel would crawl through it
    c = host.gen_powerlaw(20000, 2.2, 48, 1, 3000, 4)
    check(handle, c, x_of(c), "d")
    assert handle.info("SPMV_ALGO_USED") == ALGO["tile"] and handle.info("SPMV_MAX_ROW") >= 3000


PIPE_MATS = {
    "stencil27_64": lambda: host.gen_stencil27(64, seed=2),            # banded: x prefix grows chunk by chunk
    "lap2d_600": lambda: host.gen_laplace2d(600),
    "powerlaw_300k": lambda: host.gen_powerlaw(300000, 2.2, 64, 3, 50000, 4),  # first rows reference all of x
    "stencil27_64_base1": lambda: (lambda a: host.CSR(a.m, a.n, a.nnz, a.rowptr + 1, a.colidx + 1, a.values, 1))(host.gen_stencil27(64, seed=2)),
}


@pytest.mark.parametrize("name", sorted(PIPE_MATS))
@pytest.mark.parametrize("chunks", [0, 1, 3, 32])
@pytest.mark.parametrize("direct", [0, 1])
def test_pipelined_host_product(handle, name, chunks, direct):
    """rsg_matrix_spmv_host with beta == 0 overlaps x H2D, the product and y D2H over row chunks (HOST_CHUNKS):
    every chunking must reproduce the oracle, from pageable and from page-locked caller buffers."""
    a = PIPE_MATS[name]()
    x = orc.dlarnv(a.n) - 0.5
    handle.set_option("HOST_CHUNKS", chunks)
    handle.set_option("HOST_Y_DIRECT", direct)  # 1: kernels store y into the page-locked host buffer; 0: D2H copies
    try:
        for p in ("d", "s", "m"):
            xt = np.float32 if p == "s" else np.float64
            ref = orc.spmv(a, x, p).astype(np.float64)
            tol = 1e-12 if p != "s" else 2e-5
            mat = handle.matrix_upload(a, p)
            y = np.full(a.m, np.nan, dtype=xt)
            handle.matrix_spmv_host(mat, x.astype(xt), y)
            assert np.abs(y - ref).max() <= tol * np.abs(ref).max()
            xp = torch.from_numpy(x.astype(xt)).pin_memory()
            yp = torch.full((a.m,), float("nan"), dtype=xp.dtype).pin_memory()
            for _ in range(4):  # repeated calls re-use streams / events; from the second one on the step is a CUDA graph
                yp.fill_(float("nan"))
                handle.matrix_spmv_host(mat, xp.numpy(), yp.numpy())
                assert np.abs(yp.numpy() - ref).max() <= tol * np.abs(ref).max()
            # new x in the same buffer (graph replay reads the buffer, not a snapshot), then other buffers, then alpha
            xp.mul_(-2.0)
            handle.matrix_spmv_host(mat, xp.numpy(), yp.numpy())
            assert np.abs(yp.numpy() + 2.0 * ref).max() <= 2 * tol * np.abs(ref).max()
            xq = torch.from_numpy(x.astype(xt)).pin_memory()
            yq = torch.zeros(a.m, dtype=xp.dtype).pin_memory()
            for _ in range(3):
                handle.matrix_spmv_host(mat, xq.numpy(), yq.numpy())
                assert np.abs(yq.numpy() - ref).max() <= tol * np.abs(ref).max()
            for _ in range(3):
                handle.matrix_spmv_host(mat, xq.numpy(), yq.numpy(), alpha=0.5)
                assert np.abs(yq.numpy() - 0.5 * ref).max() <= tol * np.abs(ref).max()
            handle.set_option("HOST_GRAPH", 2)
            handle.matrix_spmv_host(mat, xq.numpy(), yq.numpy())
            handle.set_option("HOST_GRAPH", 0)
            assert np.abs(yq.numpy() - ref).max() <= tol * np.abs(ref).max()
            # beta != 0 takes the one-piece path: y = 2 A x - 0.5 y0
            y0 = orc.dlarnv(a.m).astype(xt)
            yb = y0.copy()
            handle.matrix_spmv_host(mat, x.astype(xt), yb, alpha=2.0, beta=-0.5)
            refb = 2.0 * ref - 0.5 * y0.astype(np.float64)
            assert np.abs(yb - refb).max() <= 4 * tol * max(np.abs(ref).max(), 1.0)
            handle.matrix_free(mat)
    finally:
        handle.set_option("HOST_CHUNKS", 0)
        handle.set_option("HOST_Y_DIRECT", 0)


def test_row_partition_bit_exact(handle):
    for a in (host.gen_laplace2d(50), host.gen_powerlaw(30000, 2.2, 64, 2, 9000, 4), SMALL["b1_ss_base1"]()):
        A = DeviceCSR(a, "d")
        for P in (1, 2, 3, 4, 8):
            for bal in (True, False):
                assert np.array_equal(handle.row_partition(A, P, bal), orc.row_partition(a, P, bal))


# ---- BASELINE.json configs at full size: oracle comparison (the oracle takes well under a second
# per product) plus size-independent properties (linearity, A*1 = row sums).
def _config_matrix(name):
    if name == "C1":  # through a Matrix Market file, like the reference's test_spmv
        import tempfile
        a = host.gen_laplace2d(1000)
        with tempfile.TemporaryDirectory() as d:
            p = os.path.join(d, "lap1000.mtx")
            host.write_matrix_market(p, a)
            b = host.load_matrix_market(p, 0, 0)
        assert b.stored == 4996000 and np.array_equal(a.colidx, b.colidx)
        return b
    if name == "C2":
        a = host.gen_stencil27(100, seed=2)
        assert a.stored == 26463592
        return a
    if name == "C4":
        return host.gen_powerlaw(320000, 2.2, 64, 4, 100000, 4)
    raise KeyError(name)


@pytest.mark.parametrize("name", ["C1", "C2", "C4"])
def test_full_size_configs(handle, name):
    a = _config_matrix(name)
    x = orc.dlarnv(a.n) - 0.5
    yd = None
    for p in ("d", "s", "m"):
        y = check(handle, a, x, p)
        if p == "d":
            yd = y
    # C1 and C2 are stencils: the x-window kernel; C4 (hashed columns, 4 rows of 100 002 entries): tiles, the long
    # rows split into segments
    assert handle.info("SPMV_ALGO_USED") == (ALGO["tile"] if name == "C4" else ALGO["window"])
    # linearity: A(2x + 1) = 2 A x + A 1
    A = DeviceCSR(a, "d")
    xd = torch.from_numpy(x).cuda()
    y1 = handle.spmv(A, torch.ones_like(xd))
    y2 = handle.spmv(A, 2 * xd + 1)
    lin = 2 * torch.from_numpy(yd).cuda() + y1
    rs = torch.from_numpy(row_abs_sums(a, np.abs(x) * 2 + 1, "d")).cuda()
    assert bool(((y2 - lin).abs() <= 1e-13 * rs + 1e-300).all())


@pytest.mark.gpu
@pytest.mark.parametrize("precision", ["d", "s", "m"])
def test_row_range_views_of_a_resident_matrix(handle, precision):
    """DeviceCSR.view_rows: the interior / boundary pieces the multi-GPU overlap multiplies separately
    (row pointers are a slice of the parent's and keep indexing the parent's entries).  The pieces
    written into slices of one y must reproduce the oracle's product of the whole matrix."""
    a = host.gen_stencil27(20, 20, 24, seed=2)
    x = orc.dlarnv(a.n) - 0.5
    A = DeviceCSR(a, precision)
    xt = torch.float32 if precision == "s" else torch.float64
    xd = torch.from_numpy(x).cuda().to(xt)
    y = torch.full((a.m,), float("nan"), dtype=xt, device="cuda")
    cuts = [0, 401, 402, 3 * 400 + 7, a.m - 400, a.m]
    for r0, r1 in zip(cuts[:-1], cuts[1:]):
        handle.spmv(A.view_rows(r0, r1), xd, y[r0:r1])
    ref = orc.spmv(a, x, precision).astype(np.float64)
    got = y.cpu().numpy().astype(np.float64)
    tol = 1e-12 if precision == "d" else 5e-5
    assert np.abs(got - ref).max() <= tol * np.abs(ref).max()


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
def test_halo_push_kernel(handle, dtype):
    """rsg_halo_push on one device: two 'peer' buffers stand in for P2P-mapped halo slots; a NULL slot is skipped."""
    n, ns = 5000, 1237
    x = torch.from_numpy(orc.dlarnv(n)).cuda().to(dtype)
    idx = torch.from_numpy(np.random.default_rng(5).permutation(n)[:ns].astype(np.int32)).cuda()
    peers = [torch.full((ns + 8,), -7.0, dtype=dtype, device="cuda") for _ in range(3)]
    slots = [peers[0][3:].data_ptr(), 0, peers[2].data_ptr()]
    handle.halo_push(idx, x, slots)
    torch.cuda.synchronize()
    want = x[idx.long()]
    assert torch.equal(peers[0][3:3 + ns], want) and torch.equal(peers[2][:ns], want)
    assert bool((peers[1] == -7.0).all()) and bool((peers[0][:3] == -7.0).all()) and bool((peers[0][3 + ns:] == -7.0).all())


@pytest.mark.gpu
@pytest.mark.parametrize("algo", ["window", "tile"])
def test_halo_push_signal_and_in_kernel_wait(handle, algo):
    """Two ranks emulated on one device, in ONE stream (kernels that spin on each other must never be separate
    launches on one GPU): 'rank 1' pushes its boundary entries and signals, then rank 0's block (columns relabelled
    by HaloPlan) is multiplied with the wait armed -- its halo tiles acquire the flag, copy / gather the pushed
    entries, and the result must equal the oracle's rows of the full product.  Then the fused form (the product's
    own first CTAs push and signal), the one-shot property, and a wait that can never be satisfied: it must give
    up and surface as RSG_DEPENDENCY_TIMEOUT at the next rsg_synchronize."""
    from respasol_b200.dist import HaloPlan
    from respasol_b200.gpu import RsgError
    full = host.gen_stencil27(24, 24, 40, seed=2)
    starts = np.array([0, full.m // 2, full.m])
    plans = [HaloPlan(full.rows(int(starts[r]), int(starts[r + 1])), starts, r) for r in range(2)]
    needs = [p.need for p in plans]
    for p in plans:
        p.finish(needs)
    p0, p1 = plans
    x = orc.dlarnv(full.n) - 0.5
    ref = orc.spmv(full, x, "d")[:p0.n_local]
    handle.set_option("SPMV_ALGO", ALGO[algo])
    handle.set_option("SPMV_HALO_START", p0.halo_base)
    try:
        A = DeviceCSR(p0.remapped, "d")
        handle.spmv_analysis(A)
        x_ext = torch.zeros(p0.n_ext, dtype=torch.float64, device="cuda")
        x_ext[:p0.n_local] = torch.from_numpy(x[:p0.n_local]).cuda()
        x1 = torch.from_numpy(x[p0.n_local:]).cuda()                      # rank 1's slice
        send1 = torch.from_numpy(p1.my_send.astype(np.int32)).cuda()      # what rank 1 publishes
        flags = torch.zeros(32, dtype=torch.int32, device="cuda")
        slot = x_ext[p0.halo_base + 1 * p0.width:].data_ptr()             # rank 1's slot in rank 0's halo
        other = Handle(0)
        y = torch.empty(p0.n_local, dtype=torch.float64, device="cuda")
        for epoch in (1, 2):
            x_ext[p0.halo_base:].fill_(float("nan"))
            flags[0] = epoch                                              # this rank's own signal
            other.halo_push(send1, x1, [slot], [flags[1:].data_ptr()], epoch)
            handle.spmv_halo_wait(flags, 2, epoch)
            handle.spmv(A, x_ext, y)
            handle.synchronize()
            assert handle.info("SPMV_ALGO_USED") == ALGO[algo]
            assert int(flags[1].item()) == epoch
            got = y.cpu().numpy()
            assert np.abs(got - ref).max() <= 1e-12 * np.abs(ref).max()
        # fused form: rank 0's own product launch pushes ITS boundary entries into 'rank 1's' halo and signals
        send0 = torch.from_numpy(p0.my_send.astype(np.int32)).cuda()
        peer_halo = torch.full((p1.n_ext,), float("nan"), dtype=torch.float64, device="cuda")
        peer_flags = torch.zeros(32, dtype=torch.int32, device="cuda")
        flags[:2] = 7
        torch.cuda.synchronize()
        handle.spmv_halo_fuse(send0, [0, peer_halo[p1.halo_base + 0 * p1.width:].data_ptr()], [0, peer_flags.data_ptr()], 7)
        handle.spmv_halo_wait(flags, 2, 7)
        y.fill_(0)
        handle.spmv(A, x_ext, y)
        handle.synchronize()
        assert int(peer_flags[0].item()) == 7
        want = torch.from_numpy(x[:p0.n_local][p0.my_send]).cuda()
        assert torch.equal(peer_halo[p1.halo_base:p1.halo_base + len(p0.my_send)], want)
        assert np.abs(y.cpu().numpy() - ref).max() <= 1e-12 * np.abs(ref).max()
        # one-shot: the next plain product neither pushes nor waits
        peer_flags.zero_()
        handle.spmv(A, x_ext, y)
        handle.synchronize()
        assert int(peer_flags[0].item()) == 0 and np.abs(y.cpu().numpy() - ref).max() <= 1e-12 * np.abs(ref).max()
        # a flag that never arrives: the wait gives up after RSG_OPT_WAIT_SPINS polls and the handle reports it once
        handle.set_option("WAIT_SPINS", 3000)
        handle.spmv_halo_wait(flags, 2, 1000)
        handle.spmv(A, x_ext, y)
        with pytest.raises(RsgError) as ei:
            handle.synchronize()
        assert ei.value.status == 7  # RSG_DEPENDENCY_TIMEOUT
        handle.synchronize()
        other.close()
    finally:
        handle.set_option("WAIT_SPINS", 0)
        handle.set_option("SPMV_ALGO", 0)
        handle.set_option("SPMV_HALO_START", 0)


@pytest.mark.gpu
@pytest.mark.parametrize("algo", ["window", "tile", "vector", "merge"])
@pytest.mark.parametrize("precision", ["d", "s", "m"])
def test_guard_bands(handle, algo, precision):
    """compute-sanitizer is closed on this pool (profiles/r2_sanitizer_closed.txt): the arrays of a product are embedded in
    larger buffers whose guard bands hold NaN values, wild column indices and canaries.  A read past an array's end poisons
    the result (or faults), a write outside y kills a canary.  Offsets keep the 16-byte alignment the bulk copies need."""
    G = 64  # guard elements on each side
    for a in (host.gen_stencil27(13, 11, 9, seed=3), host.gen_laplace2d(31), host.gen_powerlaw(4000, 2.2, 40, 2, 3100, 4)):
        vt = torch.float64 if precision == "d" else torch.float32
        xt = torch.float32 if precision == "s" else torch.float64
        A = DeviceCSR(a, precision)
        rp = torch.full((a.m + 1 + 2 * G,), 2 ** 30, dtype=torch.int32, device="cuda")
        ci = torch.full((a.stored + 2 * G,), 2 ** 30, dtype=torch.int32, device="cuda")
        va = torch.full((a.stored + 2 * G,), float("nan"), dtype=vt, device="cuda")
        xg = torch.full((a.n + 2 * G,), float("nan"), dtype=xt, device="cuda")
        yg = torch.full((a.m + 2 * G,), -77.0, dtype=xt, device="cuda")
        rp[G:G + a.m + 1] = A.rowptr; ci[G:G + a.stored] = A.colidx; va[G:G + a.stored] = A.values
        x = orc.dlarnv(a.n) - 0.5
        xg[G:G + a.n] = torch.from_numpy(x).cuda().to(xt)
        A.rowptr, A.colidx, A.values = rp[G:G + a.m + 1], ci[G:G + a.stored], va[G:G + a.stored]
        handle.set_option("SPMV_ALGO", ALGO[algo])
        try:
            handle.spmv(A, xg[G:G + a.n], yg[G:G + a.m])
            handle.synchronize()
        finally:
            handle.set_option("SPMV_ALGO", 0)
        assert bool((yg[:G] == -77.0).all()) and bool((yg[G + a.m:] == -77.0).all()), "write outside y"
        got = yg[G:G + a.m].cpu().numpy().astype(np.float64)
        ref = orc.spmv(a, x, precision).astype(np.float64)
        assert np.isfinite(got).all(), "a guard-band NaN reached the result"
        assert np.abs(got - ref).max() <= (1e-12 if precision == "d" else 5e-5) * np.abs(ref).max()
