// rsg_loader.cpp -- Matrix Market (coordinate) -> CSR, a from-scratch replacement for the
// reference's ReadMatrixMarket library that reproduces its OUTPUT bit for bit
// (/root/reference/ReadMatrixMarket/loadMatrixMarket.cpp:47-253 + mm_io.cpp:54-158,357-468),
// pinned by tests/test_loader.py against the reference loader compiled into oracle/_ref.
//
// Design differences (the reference tokenises with one fscanf per entry, cpp:123 / mm_io.cpp:382):
// the whole file is read once and tokenised in memory with strtol/strtod -- the same glibc
// conversions scanf uses, so parsed doubles are bit-identical -- and rowptr is allocated m+1
// entries (the reference allocates `count`).
#include "rsg_host.h"

#include <algorithm>
#include <cctype>
#include <cerrno>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

namespace {

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

} // namespace

extern "C" int rsg_load_matrix_market(const char *file, CSR *matrix, int outputBase, int transpose, int flags)
{
    if (!file || !matrix) return 0;
    matrix->rowptr = nullptr; matrix->colidx = nullptr; matrix->values = nullptr;
    matrix->m = matrix->n = matrix->nnz = 0; matrix->isSymmetric = 0;

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
    for (;;) {
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
        std::vector<int> fill(rowptr, rowptr + m);
        for (size_t l = 0; l < stored; ++l) {
            const int dst = fill[ci[l] - 1]++;
            matrix->values[dst] = cv[l];
            matrix->colidx[dst] = cj[l] - 1 + outputBase;
        }
    }
#pragma omp parallel for schedule(dynamic, 1024)
    for (int i = 0; i < m; ++i) {
        const int lo = rowptr[i], hi = rowptr[i + 1] - 1;
        if (!strictly_ascending(matrix->colidx, lo, hi))
            sort_row_like_reference(matrix->colidx, matrix->values, lo, hi);
    }
    if (outputBase != 0)
        for (int i = 0; i <= m; ++i) rowptr[i] += outputBase;

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
