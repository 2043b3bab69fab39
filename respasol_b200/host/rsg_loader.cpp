// rsg_loader.cpp -- Matrix Market (coordinate) -> CSR, a from-scratch replacement for the
// reference's ReadMatrixMarket library that reproduces its OUTPUT bit for bit
// (/root/reference/ReadMatrixMarket/loadMatrixMarket.cpp:47-253 + mm_io.cpp:54-158,357-468),
// pinned by tests/test_loader.py against the reference loader compiled into oracle/_ref.
//
// Design differences (the reference tokenises with one fscanf per entry, cpp:123 / mm_io.cpp:382):
// the whole file is read once and tokenised in memory with strtol/strtod -- the same glibc
// conversions scanf uses, so parsed doubles are bit-identical -- and rowptr is allocated m+1
// entries (the reference allocates `count`).
// Ingest is parallel (the reference's single-thread fscanf loop is the wall-time bottleneck of
// every driver): the entry region is cut at line boundaries into one chunk per OpenMP thread and tokenised
// concurrently, the COO -> CSR counting sort runs over disjoint row ranges per thread (stable, so rows holding
// duplicates enter the per-row sort in file order exactly as in the reference).  Anything irregular -- an entry
// split across lines, a stray token, an out-of-range index, a count that disagrees with the header -- makes the
// parallel pass give up and the sequential tokeniser, which reproduces the reference's behaviour token by token
// including its error paths, takes over.
#include "rsg_host.h"

#include <algorithm>
#include <cctype>
#include <cerrno>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <chrono>
#include <vector>

#ifdef _OPENMP
#include <omp.h>
#endif

namespace {

// RSG_LOADER_TRACE=1: stage timings on stderr
struct Trace {
    bool on = getenv("RSG_LOADER_TRACE") != nullptr;
    std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
    void lap(const char *what)
    {
        if (!on) return;
        auto t1 = std::chrono::steady_clock::now();
        fprintf(stderr, "[rsg_loader] %-28s %8.1f ms\n", what, std::chrono::duration<double, std::milli>(t1 - t0).count());
        t0 = t1;
    }
};

template <typename T> T *alloc64(size_t len)
{
    void *p = nullptr;
    size_t bytes = sizeof(T) * (len ? len : 1);
    if (posix_memalign(&p, 64, (bytes + 63) / 64 * 64) != 0) return nullptr;
    return static_cast<T *>(p);
}

struct Cursor {
    const char *p, *end;
    void skip_ws() { while (p < end && isspace((unsigned char)*p)) ++p; }
    bool at_end() { skip_ws(); return p >= end; }
    // %d semantics: optional sign + digits
    bool read_int(long long *v)
    {
        skip_ws();
        if (p >= end) return false;
        char *q;
        errno = 0;
        long long r = strtoll(p, &q, 10);
        if (q == p) return false;
        p = q;
        *v = r;
        return true;
    }
    bool read_double(double *v)
    {
        skip_ws();
        if (p >= end) return false;
        char *q;
        double r = strtod(p, &q);
        if (q == p) return false;
        p = q;
        *v = r;
        return true;
    }
    // consume up to and including the next newline; returns the line
    std::string read_line()
    {
        const char *s = p;
        while (p < end && *p != '\n') ++p;
        std::string line(s, p);
        if (p < end) ++p;
        return line;
    }
};

void lower(std::string &s) { for (auto &c : s) c = (char)tolower((unsigned char)c); }

// The reference sorts each row with a hand-rolled quicksort (loadMatrixMarket.cpp:5-26):
// the middle element becomes the pivot, a single left-to-right sweep gathers the smaller
// keys behind it, and the two sides recurse.  It is not stable, so rows that hold duplicate
// columns come out in an order that depends on that exact scheme; we follow the same scheme
// (own formulation, explicit stack) so that even those rows match.
void sort_row_like_reference(int *key, double *val, int lo, int hi)
{
    std::vector<std::pair<int, int>> todo;
    todo.emplace_back(lo, hi);
    while (!todo.empty()) {
        auto [l, r] = todo.back();
        todo.pop_back();
        if (l >= r) continue;
        const int mid = l + (r - l) / 2;
        std::swap(key[l], key[mid]);
        std::swap(val[l], val[mid]);
        int store = l;
        for (int i = l + 1; i <= r; ++i) {
            if (key[i] < key[l]) {
                ++store;
                std::swap(key[store], key[i]);
                std::swap(val[store], val[i]);
            }
        }
        std::swap(key[l], key[store]);
        std::swap(val[l], val[store]);
        todo.emplace_back(l, store - 1);
        todo.emplace_back(store + 1, r);
    }
}

bool strictly_ascending(const int *key, int lo, int hi)
{
    for (int i = lo; i < hi; ++i)
        if (key[i] >= key[i + 1]) return false;
    return true;
}

// Parallel tokeniser of the entry region [p0, end): true on success (exactly `nnz` well-formed one-per-line entries).
// field: 'R' real, 'I' integer, 'C' complex, 'P' pattern.  x/y come out as in the file (after the transpose swap).
bool parse_entries_parallel(const char *p0, const char *end, char field, int nnz, int m, int n, int transpose,
                            std::vector<int> &ci, std::vector<int> &cj, std::vector<double> &cv, bool *zero_based)
{
#ifndef _OPENMP
    (void)p0; (void)end; (void)field; (void)nnz; (void)m; (void)n; (void)transpose; (void)ci; (void)cj; (void)cv; (void)zero_based;
    return false;
#else
    const size_t len = (size_t)(end - p0);
    int T = omp_get_max_threads();
    if (const char *e = getenv("RSG_LOADER_THREADS")) T = atoi(e);
    if (T < 2 || len < ((size_t)1 << 20)) return false;
    T = std::min(T, 64);
    std::vector<const char *> cut((size_t)T + 1);
    cut[0] = p0;
    cut[(size_t)T] = end;
    for (int t = 1; t < T; ++t) {
        const char *q = p0 + len * (size_t)t / (size_t)T;
        while (q < end && *q != '\n') ++q;
        cut[(size_t)t] = q < end ? q + 1 : end;
    }
    // pass 1: non-blank lines per chunk = entries per chunk when the file is regular; prefix sums give every
    // thread its slice of the output arrays, so pass 2 writes in place (no per-thread buffers, no concatenation)
    std::vector<size_t> off((size_t)T + 1, 0);
#pragma omp parallel for schedule(static, 1) num_threads(T)
    for (int t = 0; t < T; ++t) {
        const char *p = cut[(size_t)t], *pe = cut[(size_t)t + 1];
        size_t lines = 0;
        while (p < pe) {
            const char *nl = static_cast<const char *>(memchr(p, '\n', (size_t)(pe - p)));
            const char *le = nl ? nl : pe;
            bool blank = true;
            for (const char *q = p; q < le; ++q)
                if (*q != ' ' && *q != '\t' && *q != '\r') { blank = false; break; }
            lines += blank ? 0 : 1;
            p = le + 1;
        }
        off[(size_t)t + 1] = lines;
    }
    for (int t = 0; t < T; ++t) off[(size_t)t + 1] += off[(size_t)t];
    if (off[(size_t)T] != (size_t)nnz) return false;
    ci.resize((size_t)nnz); cj.resize((size_t)nnz); cv.resize((size_t)nnz);
    const bool pattern = field == 'P';
    int bad_any = 0, zero_any = 0;
#pragma omp parallel for schedule(static, 1) num_threads(T) reduction(| : bad_any, zero_any)
    for (int t = 0; t < T; ++t) {
        const char *p = cut[(size_t)t], *pe = cut[(size_t)t + 1];
        size_t k = off[(size_t)t];
        const size_t kend = off[(size_t)t + 1];
        bool bad = false, zero = false;
        auto blanks = [&](void) { while (p < pe && (*p == ' ' || *p == '\t' || *p == '\r')) ++p; };
        while (p < pe && !bad) {
            blanks();
            if (p >= pe) break;
            if (*p == '\n') { ++p; continue; }
            char *q;
            long long x = strtoll(p, &q, 10);
            if (q == p) { bad = true; break; }
            p = q; blanks();
            if (p >= pe || *p == '\n') { bad = true; break; }
            long long y = strtoll(p, &q, 10);
            if (q == p) { bad = true; break; }
            p = q;
            double re = 1.0;
            if (!pattern) {
                blanks();
                if (p >= pe || *p == '\n') { bad = true; break; }
                if (field == 'I') { long long iv = strtoll(p, &q, 10); re = (double)iv; }
                else re = strtod(p, &q);
                if (q == p) { bad = true; break; }
                p = q;
                if (field == 'C') {
                    blanks();
                    if (p >= pe || *p == '\n') { bad = true; break; }
                    (void)strtod(p, &q);
                    if (q == p) { bad = true; break; }
                    p = q;
                }
            }
            blanks();
            if (p < pe && *p != '\n') { bad = true; break; } // a stray token: not one entry per line
            if (transpose) std::swap(x, y);
            if (x < 0 || y < 0 || x > m || y > n || k >= kend) { bad = true; break; }
            if (x == 0 || y == 0) zero = true;
            ci[k] = (int)x; cj[k] = (int)y; cv[k] = re;
            ++k;
        }
        if (k != kend) bad = true;
        bad_any |= bad ? 1 : 0;
        zero_any |= zero ? 1 : 0;
    }
    if (bad_any) { ci.clear(); cj.clear(); cv.clear(); return false; }
    const bool zero = zero_any != 0;
    *zero_based = zero;
    return true;
#endif
}

} // namespace

extern "C" int rsg_load_matrix_market(const char *file, CSR *matrix, int outputBase, int transpose, int flags)
{
    if (!file || !matrix) return 0;
    matrix->rowptr = nullptr; matrix->colidx = nullptr; matrix->values = nullptr;
    matrix->m = matrix->n = matrix->nnz = 0; matrix->isSymmetric = 0;

    Trace trace;
    FILE *fp = fopen(file, "rb");
    if (!fp) {
        // the reference exit(-1)s here (cpp:50-53); a library must not
        fprintf(stderr, "Failed to open file %s\n", file);
        return 0;
    }
    std::vector<char> buf;
    {
        fseek(fp, 0, SEEK_END);
        long sz = ftell(fp);
        fseek(fp, 0, SEEK_SET);
        buf.resize((size_t)(sz > 0 ? sz : 0) + 1);
        size_t got = sz > 0 ? fread(buf.data(), 1, (size_t)sz, fp) : 0;
        buf[got] = '\0';
        buf.resize(got + 1);
        fclose(fp);
    }
    Cursor cur{buf.data(), buf.data() + buf.size() - 1};
    trace.lap("read file");

    // ---- banner: five whitespace-separated tokens on line 1 (mm_io.cpp:54-158)
    char kind = 0, field = 0, storage = 0;
    {
        std::string line = cur.read_line();
        char t[5][65];
        if (sscanf(line.c_str(), "%64s %64s %64s %64s %64s", t[0], t[1], t[2], t[3], t[4]) != 5 ||
            strncmp(t[0], "%%MatrixMarket", 14) != 0) {
            fprintf(stderr, "Error: could not process Matrix Market banner.\n");
            return 0;
        }
        std::string obj(t[1]), fmt(t[2]), fld(t[3]), sto(t[4]);
        lower(obj); lower(fmt); lower(fld); lower(sto);
        bool ok = (obj == "matrix");
        if (fmt == "coordinate") kind = 'C'; else if (fmt == "array") kind = 'A'; else ok = false;
        if (fld == "real") field = 'R'; else if (fld == "complex") field = 'C';
        else if (fld == "pattern") field = 'P'; else if (fld == "integer") field = 'I'; else ok = false;
        if (sto == "general") storage = 'G'; else if (sto == "symmetric") storage = 'S';
        else if (sto == "hermitian") storage = 'H'; else if (sto == "skew-symmetric") storage = 'K';
        else ok = false;
        if (!ok) {
            fprintf(stderr, "Error: could not process Matrix Market banner.\n");
            return 0;
        }
    }
    // validity rules of mm_is_valid (mm_io.cpp:28-50) + the loader's own rejection of arrays (cpp:63)
    if (kind == 'A' || (field == 'R' && storage == 'H') ||
        (field == 'P' && (storage == 'H' || storage == 'K'))) {
        fprintf(stderr, "Error: only support sparse and real matrices.\n");
        return 0;
    }
    const bool pattern = (field == 'P');

    // ---- size line: skip '%' lines, then three integers (mm_io.cpp:404-468)
    long long M = 0, N = 0, NZ = 0;
    for (;;) {
        if (cur.p >= cur.end) { fprintf(stderr, "Error: could not read matrix size.\n"); return 0; }
        if (*cur.p == '%') { cur.read_line(); continue; }
        break;
    }
    if (!cur.read_int(&M) || !cur.read_int(&N) || !cur.read_int(&NZ) || M < 0 || N < 0 || NZ < 0 ||
        M > INT32_MAX || N > INT32_MAX || NZ > INT32_MAX) {
        fprintf(stderr, "Error: could not read matrix size.\n");
        return 0;
    }
    int m = (int)M, n = (int)N;
    const int nnz = (int)NZ;
    if (transpose) std::swap(m, n);
    const bool symmetric = (storage == 'S'); // only 'S' expands
    matrix->isSymmetric = symmetric ? 1 : 0;

    // ---- entries (loadMatrixMarket.cpp:117-160)
    std::vector<int> ci, cj;
    std::vector<double> cv;
    ci.reserve((size_t)nnz); cj.reserve((size_t)nnz); cv.reserve((size_t)nnz);
    bool zero_based_file = false;
    const bool parsed = parse_entries_parallel(cur.p, cur.end, field, nnz, m, n, transpose, ci, cj, cv, &zero_based_file);
    while (!parsed) { // sequential tokeniser: small files, irregular files, and every error path
        long long x, y;
        double re = 0, im = 0;
        const char *save = cur.p;
        if (!cur.read_int(&x) || !cur.read_int(&y)) { cur.p = save; break; }
        if (field == 'R') { if (!cur.read_double(&re)) break; }
        else if (field == 'I') { long long iv; if (!cur.read_int(&iv)) break; re = (double)iv; }
        else if (field == 'C') { if (!cur.read_double(&re) || !cur.read_double(&im)) break; }
        if (transpose) std::swap(x, y);
        if (x > m || y > n) { // checked before any 0-based shift, like cpp:126
            fprintf(stderr, "Error: (%lld %lld) coordinate is out of range.\n", x, y);
            return 0;
        }
        if ((long long)ci.size() >= (long long)nnz + 1) break; // already too many; reported below
        ci.push_back((int)x); cj.push_back((int)y); cv.push_back(pattern ? 1.0 : re);
        if (x == 0 || y == 0) zero_based_file = true;
    }
    trace.lap(parsed ? "tokenise (parallel)" : "tokenise (sequential)");
    if (zero_based_file)
        for (size_t i = 0; i < ci.size(); ++i) { ci[i]++; cj[i]++; }
    if ((long long)ci.size() != nnz) {
        fprintf(stderr, "Error: nnz (%d) specified in the header doesn't match with # of lines (%zu) in file %s\n",
                nnz, ci.size(), file);
        return 0;
    }
    for (int i = 0; i < nnz; ++i)
        if (ci[i] < 1 || cj[i] < 1 || ci[i] > m || cj[i] > n) {
            fprintf(stderr, "Error: (%d %d) coordinate is out of range.\n", ci[i], cj[i]);
            return 0;
        }

    // ---- symmetric: count (and in fixed mode, append) mirrors that are not already present (cpp:162-200)
    size_t count = (size_t)nnz;
    if (symmetric) {
        std::vector<int> rp((size_t)m + 1, 0), cols((size_t)nnz);
        for (int i = 0; i < nnz; ++i) rp[ci[i]]++;
        for (int i = 0; i < m; ++i) rp[i + 1] += rp[i];
        {
            std::vector<int> fill(rp.begin(), rp.end() - 1);
            for (int i = 0; i < nnz; ++i) cols[fill[ci[i] - 1]++] = cj[i];
        }
        for (int i = 0; i < m; ++i) std::sort(cols.begin() + rp[i], cols.begin() + rp[i + 1]);
        for (int i = 0; i < nnz; ++i) {
            const int x = ci[i], y = cj[i];
            if (x == y) continue;
            const bool present = (y <= m) && std::binary_search(cols.begin() + rp[y - 1], cols.begin() + rp[y], x);
            if (!present) {
                ++count;
                if (flags & RSG_LOADER_EXPAND_SYMMETRIC) { ci.push_back(y); cj.push_back(x); cv.push_back(cv[i]); }
            }
        }
    }
    // entries that really go into the CSR: the reference converts only the first `nnz` (cpp:220,225)
    const size_t stored = (flags & RSG_LOADER_EXPAND_SYMMETRIC) ? ci.size() : (size_t)nnz;
    if ((flags & RSG_LOADER_EXPAND_SYMMETRIC) && symmetric)
        for (size_t i = (size_t)nnz; i < stored; ++i)
            if (ci[i] > m || cj[i] > n) { fprintf(stderr, "Error: symmetric file is not square.\n"); return 0; }

    trace.lap("checks / symmetric count");
    // ---- COO -> CSR: stable counting sort by row, then per-row sort by column (cpp:202-242)
    matrix->values = alloc64<double>(std::max(count, stored));
    matrix->colidx = alloc64<int>(std::max(count, stored));
    matrix->rowptr = alloc64<int>((size_t)m + 1);
    if (!matrix->values || !matrix->colidx || !matrix->rowptr) {
        fprintf(stderr, "Failed to allocate memory\n");
        rsg_free_csr(matrix);
        return 0;
    }
    int *rowptr = matrix->rowptr;
    for (int i = 0; i <= m; ++i) rowptr[i] = 0;
    for (size_t l = 0; l < stored; ++l) rowptr[ci[l]]++;
    for (int i = 0; i < m; ++i) rowptr[i + 1] += rowptr[i];
    {
        // stable scatter: every thread owns a contiguous range of rows and walks the whole entry list for them,
        // so entries of a row land in file order whatever the thread count
        std::vector<int> fill(rowptr, rowptr + m);
        int T = 1;
#ifdef _OPENMP
        T = stored >= ((size_t)1 << 18) ? std::min(omp_get_max_threads(), 64) : 1;
#endif
#pragma omp parallel for schedule(static, 1) num_threads(T)
        for (int t = 0; t < T; ++t) {
            const int r0 = (int)((long long)m * t / T) + 1, r1 = (int)((long long)m * (t + 1) / T); // 1-based rows [r0, r1]
            for (size_t l = 0; l < stored; ++l) {
                const int r = ci[l];
                if (r < r0 || r > r1) continue;
                const int dst = fill[r - 1]++;
                matrix->values[dst] = cv[l];
                matrix->colidx[dst] = cj[l] - 1 + outputBase;
            }
        }
    }
    trace.lap("counting sort by row");
#pragma omp parallel for schedule(dynamic, 1024)
    for (int i = 0; i < m; ++i) {
        const int lo = rowptr[i], hi = rowptr[i + 1] - 1;
        if (!strictly_ascending(matrix->colidx, lo, hi))
            sort_row_like_reference(matrix->colidx, matrix->values, lo, hi);
    }
    if (outputBase != 0)
        for (int i = 0; i <= m; ++i) rowptr[i] += outputBase;
    trace.lap("per-row column sort");

    matrix->m = m;
    matrix->n = n;
    matrix->nnz = (int)count; // the reference reports the expanded count even though it stored nnz (9-D1)
    return 1;
}

extern "C" int rsg_csr_stored_nnz(const CSR *a)
{
    return (a && a->rowptr) ? a->rowptr[a->m] - a->rowptr[0] : 0;
}

extern "C" void rsg_free_csr(CSR *a)
{
    if (!a) return;
    free(a->rowptr); free(a->colidx); free(a->values);
    a->rowptr = nullptr; a->colidx = nullptr; a->values = nullptr;
}

extern "C" int rsg_write_matrix_market(const char *file, const CSR *a, int base)
{
    FILE *fp = fopen(file, "w");
    if (!fp) return 0;
    const int stored = a->rowptr[a->m] - a->rowptr[0];
    fprintf(fp, "%%%%MatrixMarket matrix coordinate real general\n%d %d %d\n", a->m, a->n, stored);
    std::vector<char> out;
    out.reserve(1 << 22);
    char line[96];
    for (int i = 0; i < a->m; ++i) {
        for (int k = a->rowptr[i] - base; k < a->rowptr[i + 1] - base; ++k) {
            int len = snprintf(line, sizeof line, "%d %d %.17g\n", i + 1, a->colidx[k] - base + 1, a->values[k]);
            out.insert(out.end(), line, line + len);
        }
        if (out.size() > (1u << 22) - 4096) { fwrite(out.data(), 1, out.size(), fp); out.clear(); }
    }
    fwrite(out.data(), 1, out.size(), fp);
    return fclose(fp) == 0 ? 1 : 0;
}
