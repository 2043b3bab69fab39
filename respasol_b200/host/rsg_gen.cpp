// rsg_gen.cpp -- deterministic synthetic matrices shaped like the SuiteSparse groups the
// reference benchmarks (/root/reference/matrices/*/get*.sh need the network; README.md:28-31
// gives the size classes).  Shapes, seeds and sizes follow SURVEY.md 8d / BASELINE.json configs.
// PRNG: splitmix64, used as a stateless hash so a value depends only on (seed,row,col) and a
// row slab generated on one GPU equals the same rows of the full matrix.
#include "rsg_host.h"

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <vector>

namespace {

inline uint64_t splitmix64(uint64_t x)
{
    x += 0x9E3779B97F4A7C15ULL;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ULL;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBULL;
    return x ^ (x >> 31);
}
inline uint64_t hash3(uint64_t seed, uint64_t a, uint64_t b)
{
    return splitmix64(splitmix64(splitmix64(seed) ^ a) ^ (b * 0xD6E8FEB86659FD93ULL));
}
inline double u01(uint64_t h) { return (double)(h >> 11) * (1.0 / 9007199254740992.0); }

template <typename T> T *alloc64(size_t len)
{
    void *p = nullptr;
    size_t bytes = sizeof(T) * (len ? len : 1);
    if (posix_memalign(&p, 64, (bytes + 63) / 64 * 64) != 0) return nullptr;
    return static_cast<T *>(p);
}

bool alloc_csr(CSR *out, int m, int n, size_t nnz)
{
    out->isSymmetric = 0;
    out->m = m; out->n = n; out->nnz = (int)nnz;
    out->rowptr = alloc64<int>((size_t)m + 1);
    out->colidx = alloc64<int>(nnz);
    out->values = alloc64<double>(nnz);
    if (!out->rowptr || !out->colidx || !out->values) { rsg_free_csr(out); return false; }
    return true;
}

} // namespace

extern "C" void rsg_dlarnv_u01(int64_t n, double *x)
{
    const uint64_t mult = 33952834046453ULL, mask = (1ULL << 48) - 1;
    uint64_t s = 1;
    for (int64_t i = 0; i < n; ++i) {
        s = (s * mult) & mask;
        x[i] = (double)s * (1.0 / 281474976710656.0);
    }
}

extern "C" int rsg_gen_laplace2d(int N, CSR *out)
{
    if (N < 1 || !out) return 0;
    const int64_t m64 = (int64_t)N * N, nnz64 = 5 * m64 - 4 * (int64_t)N;
    if (m64 > INT32_MAX || nnz64 > INT32_MAX) return 0;
    if (!alloc_csr(out, (int)m64, (int)m64, (size_t)nnz64)) return 0;
    int k = 0;
    for (int i = 0; i < N; ++i)
        for (int j = 0; j < N; ++j) {
            const int r = i * N + j;
            out->rowptr[r] = k;
            if (i > 0)     { out->colidx[k] = r - N; out->values[k++] = -1.0; }
            if (j > 0)     { out->colidx[k] = r - 1; out->values[k++] = -1.0; }
                             out->colidx[k] = r;     out->values[k++] = 4.0;
            if (j < N - 1) { out->colidx[k] = r + 1; out->values[k++] = -1.0; }
            if (i < N - 1) { out->colidx[k] = r + N; out->values[k++] = -1.0; }
        }
    out->rowptr[m64] = k;
    return 1;
}

extern "C" int rsg_gen_stencil27(int nx, int ny, int nz, int z0, int z1, uint64_t seed, CSR *out)
{
    if (nx < 1 || ny < 1 || nz < 1 || z0 < 0 || z1 > nz || z0 >= z1 || !out) return 0;
    const int64_t n64 = (int64_t)nx * ny * nz, m64 = (int64_t)nx * ny * (z1 - z0);
    if (n64 > INT32_MAX) return 0;
    auto span = [](int c, int lim) { return (c > 0) + 1 + (c < lim - 1); };
    // exact nnz of the slab: sum over planes of (3nx-2)(3ny-2)*zspan
    int64_t nnz64 = 0;
    for (int z = z0; z < z1; ++z) nnz64 += (int64_t)(3 * nx - 2) * (3 * ny - 2) * span(z, nz);
    if (nnz64 > INT32_MAX) return 0;
    if (!alloc_csr(out, (int)m64, (int)n64, (size_t)nnz64)) return 0;
    // row starts first (closed form per row), then fill rows in parallel
    {
        int64_t k = 0;
        int r = 0;
        for (int z = z0; z < z1; ++z)
            for (int y = 0; y < ny; ++y)
                for (int x = 0; x < nx; ++x) {
                    out->rowptr[r++] = (int)k;
                    k += span(x, nx) * span(y, ny) * span(z, nz);
                }
        out->rowptr[r] = (int)k;
    }
#pragma omp parallel for schedule(static)
    for (int zz = z0; zz < z1; ++zz)
        for (int y = 0; y < ny; ++y)
            for (int x = 0; x < nx; ++x) {
                const int lr = ((zz - z0) * ny + y) * nx + x;
                const int64_t gr = ((int64_t)zz * ny + y) * nx + x;
                int k = out->rowptr[lr];
                for (int dz = -1; dz <= 1; ++dz) {
                    if (zz + dz < 0 || zz + dz >= nz) continue;
                    for (int dy = -1; dy <= 1; ++dy) {
                        if (y + dy < 0 || y + dy >= ny) continue;
                        for (int dx = -1; dx <= 1; ++dx) {
                            if (x + dx < 0 || x + dx >= nx) continue;
                            const int64_t gc = ((int64_t)(zz + dz) * ny + (y + dy)) * nx + (x + dx);
                            out->colidx[k] = (int)gc;
                            if (gc == gr) out->values[k] = 26.0;
                            else if (seed == 0) out->values[k] = -1.0;
                            else out->values[k] = -(1.0 + 0.01 * (2.0 * u01(hash3(seed, (uint64_t)gr, (uint64_t)gc)) - 1.0));
                            ++k;
                        }
                    }
                }
            }
    return 1;
}

extern "C" int rsg_gen_femlike(int N, double p, uint64_t seed, CSR *out)
{
    if (N < 1 || !out) return 0;
    const int64_t n64 = (int64_t)N * N * N;
    if (n64 > INT32_MAX / 27) return 0;
    const int n = (int)n64;
    std::vector<int> rp((size_t)n + 1, 0);
    std::vector<int> cols;
    std::vector<double> vals;
    cols.reserve((size_t)n * 18);
    vals.reserve((size_t)n * 18);
    for (int z = 0; z < N; ++z)
        for (int y = 0; y < N; ++y)
            for (int x = 0; x < N; ++x) {
                const int r = (z * N + y) * N + x;
                const size_t row_begin = cols.size();
                size_t diag_at = 0;
                double offsum = 0.0;
                for (int dz = -1; dz <= 1; ++dz)
                    for (int dy = -1; dy <= 1; ++dy)
                        for (int dx = -1; dx <= 1; ++dx) {
                            const int xx = x + dx, yy = y + dy, zz = z + dz;
                            if (xx < 0 || xx >= N || yy < 0 || yy >= N || zz < 0 || zz >= N) continue;
                            const int c = (zz * N + yy) * N + xx;
                            if (c == r) { diag_at = cols.size(); cols.push_back(c); vals.push_back(0.0); continue; }
                            const uint64_t lo = (uint64_t)std::min(r, c), hi = (uint64_t)std::max(r, c);
                            if (u01(hash3(seed, lo, hi)) >= p) continue; // the PAIR (r,c),(c,r) is dropped together
                            const double v = -(0.5 + u01(hash3(seed ^ 0xA5A5A5A5ULL, (uint64_t)r, (uint64_t)c)));
                            cols.push_back(c); vals.push_back(v);
                            offsum += std::fabs(v);
                        }
                vals[diag_at] = offsum > 0.0 ? 1.05 * offsum : 1.0;
                (void)row_begin;
                rp[(size_t)r + 1] = (int)cols.size();
            }
    if (!alloc_csr(out, n, n, cols.size())) return 0;
    memcpy(out->rowptr, rp.data(), sizeof(int) * ((size_t)n + 1));
    memcpy(out->colidx, cols.data(), sizeof(int) * cols.size());
    memcpy(out->values, vals.data(), sizeof(double) * vals.size());
    return 1;
}

extern "C" int rsg_gen_powerlaw(int m, double alpha, int maxlen, int ndense, int dense_len, uint64_t seed, CSR *out)
{
    if (m < 1 || maxlen < 1 || ndense < 0 || dense_len < 0 || dense_len > m || !out) return 0;
    std::vector<std::vector<int>> rows((size_t)m);
    // sparse part: Pareto-distributed lengths, random distinct columns, plus the diagonal
    for (int i = 0; i < m; ++i) {
        const double u = 1.0 - u01(hash3(seed, (uint64_t)i, 0x51ED)); // (0,1]
        int len = 1 + (int)std::floor(std::pow(u, -1.0 / alpha));        // 1 + floor(Pareto(xm=1))
        len = std::min(std::max(len, 1), maxlen);
        auto &r = rows[(size_t)i];
        r.push_back(i);
        for (int t = 0; t < len; ++t) r.push_back((int)(hash3(seed + 1, (uint64_t)i, (uint64_t)t) % (uint64_t)m));
    }
    // dense rows at fixed, spread-out indices, and their transposed columns
    for (int d = 0; d < ndense; ++d) {
        const int dr = (int)(((int64_t)(2 * d + 1) * m) / (2 * (ndense > 0 ? ndense : 1)));
        for (int j = 0; j < dense_len; ++j) {
            const int c = (int)(((int64_t)j * m) / dense_len);
            rows[(size_t)dr].push_back(c);
            rows[(size_t)c].push_back(dr);
        }
    }
    size_t nnz = 0;
    for (auto &r : rows) {
        std::sort(r.begin(), r.end());
        r.erase(std::unique(r.begin(), r.end()), r.end());
        nnz += r.size();
    }
    if (nnz > (size_t)INT32_MAX) return 0;
    if (!alloc_csr(out, m, m, nnz)) return 0;
    int k = 0;
    for (int i = 0; i < m; ++i) {
        out->rowptr[i] = k;
        double rowsum = 0.0;
        int diag_at = -1;
        for (int c : rows[(size_t)i]) {
            out->colidx[k] = c;
            if (c == i) { diag_at = k; out->values[k] = 0.0; }
            else {
                const double v = 2.0 * u01(hash3(seed + 2, (uint64_t)i, (uint64_t)c)) - 1.0;
                out->values[k] = v;
                rowsum += std::fabs(v);
            }
            ++k;
        }
        out->values[diag_at] = rowsum + 1.0;
    }
    out->rowptr[m] = k;
    return 1;
}

// Generic shape class for the sweep harness: m rows of `per_row` entries each -- the diagonal plus per_row-1
// distinct columns at hashed offsets within +-half_band (clipped to the matrix; rows near the border are shorter
// when the band runs out of distinct columns).  Values -(0.5+u), diagonal 1.05 * sum|offdiag|.
extern "C" int rsg_gen_banded(int m, int per_row, int half_band, uint64_t seed, CSR *out)
{
    if (!out || m < 1 || per_row < 1 || half_band < 1) return 0;
    half_band = std::min(half_band, m);
    std::vector<int> rp((size_t)m + 1, 0);
    std::vector<std::vector<int>> rows;
    // two passes would double the hashing; rows are small, keep them
    std::vector<int> ci;
    ci.reserve((size_t)m * (size_t)per_row);
    std::vector<int> cur;
    for (int i = 0; i < m; ++i) {
        cur.clear();
        cur.push_back(i);
        const int lo = std::max(0, i - half_band), hi = std::min(m - 1, i + half_band);
        const int span = hi - lo + 1;
        const int want = std::min(per_row, span);
        for (uint64_t t = 0; (int)cur.size() < want && t < (uint64_t)per_row * 8; ++t) {
            const int c = lo + (int)(hash3(seed, (uint64_t)i, t) % (uint64_t)span);
            if (std::find(cur.begin(), cur.end(), c) == cur.end()) cur.push_back(c);
        }
        std::sort(cur.begin(), cur.end());
        ci.insert(ci.end(), cur.begin(), cur.end());
        rp[(size_t)i + 1] = (int)ci.size();
    }
    if (!alloc_csr(out, m, m, ci.size())) return 0;
    std::memcpy(out->rowptr, rp.data(), sizeof(int) * ((size_t)m + 1));
    std::memcpy(out->colidx, ci.data(), sizeof(int) * ci.size());
    for (int i = 0; i < m; ++i) {
        double sum = 0;
        int dk = -1;
        for (int k = rp[(size_t)i]; k < rp[(size_t)i + 1]; ++k) {
            if (ci[(size_t)k] == i) { dk = k; continue; }
            const double v = -(0.5 + u01(hash3(seed ^ 0x51ED270B, (uint64_t)i, (uint64_t)ci[(size_t)k])));
            out->values[k] = v;
            sum += std::fabs(v);
        }
        out->values[dk] = 1.05 * sum + 1e-3;
    }
    return 1;
}
