/*
 * oracle.c -- CPU restatement of the ReSpaSol SpMV / ILU0 / SpTRSV hot path.
 *
 * THIS IS TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs may load it.  The product
 * (libreaspasol_gpu.so) never links, loads or calls anything in this directory.
 *
 * Parity status (see DESIGN.md "Oracle"):
 *   - Matrix Market -> CSR: PINNED against the reference loader itself, compiled
 *     from /root/reference/ReadMatrixMarket into oracle/_ref/ (oracle/Makefile).
 *   - dlarnv(1,{0,0,0,1}) vector: PINNED (bit-identical to LAPACK dlaruv).
 *   - SpMV / ILU0 / SpTRSV arithmetic: the reference delegates these to closed
 *     third-party libraries that are absent from /root/reference (cuSPARSE: un-pinned
 *     "CUDA 10 or recent", README.md:20; MKL "2020 or recent", README.md:19; SuperLU 5.2.1,
 *     README.md:18).  The restatement follows the published algorithms and the
 *     reference's call sites; the reference holds no golden vectors for them, so
 *     for these functions PARITY IS UNPINNED beyond the KATs of SURVEY.md 10
 *     (K2 SpMV values, K4 ILU0/levels/solves), which tests/ check.
 *
 * Reference call sites restated (paths relative to /root/reference):
 *   SpMV   y = alpha*A*x + beta*y, NON_TRANSPOSE, CSR int32    GPU/spmv.cu:77-78,174-186
 *                                                               test_spmv.c:168-180
 *   ILU0   in place, natural order, no pivoting, unit L         GPU/ilu0.cu:203-209,264-268
 *   zero pivots (structural / numerical)                        GPU/ilu0.cu:221-226,278-282
 *   L solve (unit lower), U solve (non-unit upper, intended)    GPU/ilu0.cu:129-141,284-310
 *   x = dlarnv(idist=1, seed {0,0,0,1})                         test_spmv.c:75-76
 *
 * Build: gcc -O2 -ffp-contract=off -fopenmp -shared -fPIC (oracle/Makefile).
 * -ffp-contract=off: every fused multiply-add in this file is an explicit fma().
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define ORC_OK 0
#define ORC_BAD_ARG 1
#define ORC_ZERO_PIVOT_STRUCTURAL 5
#define ORC_ZERO_PIVOT_NUMERIC 6

/* ---------------------------------------------------------------- dlarnv ---
 * LAPACK dlaruv as used by LAPACKE_dlarnv(1, {0,0,0,1}, n, x), test_spmv.c:75-76:
 * 48-bit multiplicative congruential generator s <- s*33952834046453 mod 2^48,
 * u = s * 2^-48, starting from s0 = 1; output i uses s_{i+1}.            */
void orc_dlarnv_u01(int64_t n, double *x)
{
    const uint64_t mult = 33952834046453ULL, mask = (1ULL << 48) - 1;
    uint64_t s = 1;
    for (int64_t i = 0; i < n; ++i) {
        s = (s * mult) & mask;
        x[i] = (double)s * (1.0 / 281474976710656.0);
    }
}

/* ------------------------------------------------------------------ SpMV ---
 * Rows summed left to right in stored order, in the working precision
 * (SURVEY.md 8c).  beta == 0 never reads y (cuSPARSE semantics, spmv.cu:78). */
#define SPMV_BODY(TV, TX, TA)                                                   \
    const int *rp = rowptr;                                                     \
    for (int i = 0; i < m; ++i) {                                               \
        TA acc = 0;                                                             \
        for (int k = rp[i] - base; k < rp[i + 1] - base; ++k)                   \
            acc += (TA)values[k] * (TA)x[colidx[k] - base];                     \
        y[i] = (beta == 0) ? (TX)(alpha * acc) : (TX)(alpha * acc + beta * y[i]); \
    }

void orc_spmv_csr_d(int m, int base, const int *rowptr, const int *colidx,
                    const double *values, double alpha, const double *x,
                    double beta, double *y)
{
    SPMV_BODY(double, double, double)
}

void orc_spmv_csr_s(int m, int base, const int *rowptr, const int *colidx,
                    const float *values, float alpha, const float *x,
                    float beta, float *y)
{
    SPMV_BODY(float, float, float)
}

/* mixed: A stored fp32, x / y / accumulation fp64 (north_star; not in the reference) */
void orc_spmv_csr_mixed(int m, int base, const int *rowptr, const int *colidx,
                        const float *values, double alpha, const double *x,
                        double beta, double *y)
{
    SPMV_BODY(float, double, double)
}

/* Row-parallel OpenMP forms: the stand-in for mkl_sparse_{d,s}_mv
 * (test_spmv.c:168-180, bench_spmv.cc:45-48), used only as the timed CPU baseline. */
#define SPMV_OMP_BODY(TX, TA)                                                   \
    _Pragma("omp parallel for schedule(static) num_threads(nthreads)")          \
    for (int i = 0; i < m; ++i) {                                               \
        TA acc = 0;                                                             \
        for (int k = rowptr[i] - base; k < rowptr[i + 1] - base; ++k)           \
            acc += (TA)values[k] * (TA)x[colidx[k] - base];                     \
        y[i] = (beta == 0) ? (TX)(alpha * acc) : (TX)(alpha * acc + beta * y[i]); \
    }

void orc_spmv_csr_d_omp(int nthreads, int m, int base, const int *rowptr, const int *colidx,
                        const double *values, double alpha, const double *x,
                        double beta, double *y)
{
    SPMV_OMP_BODY(double, double)
}

void orc_spmv_csr_s_omp(int nthreads, int m, int base, const int *rowptr, const int *colidx,
                        const float *values, float alpha, const float *x,
                        float beta, float *y)
{
    SPMV_OMP_BODY(float, float)
}

void orc_spmv_csr_mixed_omp(int nthreads, int m, int base, const int *rowptr, const int *colidx,
                            const float *values, double alpha, const double *x,
                            double beta, double *y)
{
    SPMV_OMP_BODY(double, double)
}

int orc_max_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* ------------------------------------------------------------------ ILU0 ---
 * csrilu02 as called at GPU/ilu0.cu:264-268: IKJ, natural order, zero fill,
 * in place; L unit-lower (diagonal not stored), U upper incl. diagonal.
 *   for i: for k<i in row i ascending: a_ik /= a_kk;
 *          for j>k in row k, (i,j) in pattern: a_ij = fma(-a_ik, a_kj, a_ij)
 * Zero pivots mirror cusparseXcsrilu02_zeroPivot (ilu0.cu:222,278): structural =
 * missing diagonal entry (smallest such row), numerical = smallest row whose
 * finished pivot u_ii is exactly 0 (the factorization still runs to the end).  Columns must be sorted and unique within a row.               */
#define ILU0_BODY(T, FMA)                                                       \
    int *diag = (int *)malloc(sizeof(int) * (size_t)(n > 0 ? n : 1));           \
    int *pos = (int *)malloc(sizeof(int) * (size_t)(n > 0 ? n : 1));            \
    if (!diag || !pos) { free(diag); free(pos); return ORC_BAD_ARG; }           \
    *pivot_row = -1;                                                            \
    for (int i = 0; i < n; ++i) {                                               \
        diag[i] = -1;                                                           \
        for (int k = rowptr[i] - base; k < rowptr[i + 1] - base; ++k)           \
            if (colidx[k] - base == i) { diag[i] = k; break; }                  \
        if (diag[i] < 0) { *pivot_row = i; free(diag); free(pos);               \
                           return ORC_ZERO_PIVOT_STRUCTURAL; }                  \
    }                                                                           \
    for (int j = 0; j < n; ++j) pos[j] = -1;                                    \
    int status = ORC_OK;                                                        \
    for (int i = 0; i < n; ++i) {                                               \
        int rs = rowptr[i] - base, re = rowptr[i + 1] - base;                   \
        for (int k = rs; k < re; ++k) pos[colidx[k] - base] = k;                \
        for (int kk = rs; kk < re; ++kk) {                                      \
            int k = colidx[kk] - base;                                          \
            if (k >= i) break;                                                  \
            T lik = a[kk] / a[diag[k]];                                         \
            a[kk] = lik;                                                        \
            for (int jj = diag[k] + 1; jj < rowptr[k + 1] - base; ++jj) {       \
                int p = pos[colidx[jj] - base];                                 \
                if (p >= 0) a[p] = FMA(-lik, a[jj], a[p]);                      \
            }                                                                   \
        }                                                                       \
        for (int k = rs; k < re; ++k) pos[colidx[k] - base] = -1;               \
        if (a[diag[i]] == (T)0 && status == ORC_OK) {                           \
            status = ORC_ZERO_PIVOT_NUMERIC; *pivot_row = i; }                  \
    }                                                                           \
    free(diag); free(pos);                                                      \
    return status;

int orc_ilu0_d(int n, int base, const int *rowptr, const int *colidx, double *a, int *pivot_row)
{
    ILU0_BODY(double, fma)
}

int orc_ilu0_s(int n, int base, const int *rowptr, const int *colidx, float *a, int *pivot_row)
{
    ILU0_BODY(float, fmaf)
}

/* ---------------------------------------------------------------- SpTRSV ---
 * L z = b : L = strictly-lower part of the in-place LU values, unit diagonal
 *           (desc_L, GPU/ilu0.cu:129-134; csrsv2_solve ilu0.cu:296-298).
 * U y = z : U = upper part incl. diagonal, non-unit (desc_U fill/diag modes,
 *           ilu0.cu:136-141 -- the INTENDED second solve).
 * Sequential left-to-right sums with explicit fma.                          */
#define TRSV_L_BODY(T, FMA)                                                     \
    for (int i = 0; i < n; ++i) {                                               \
        T s = b[i];                                                             \
        for (int k = rowptr[i] - base; k < rowptr[i + 1] - base; ++k) {         \
            int j = colidx[k] - base;                                           \
            if (j >= i) break;                                                  \
            s = FMA(-lu[k], z[j], s);                                           \
        }                                                                       \
        z[i] = s;                                                               \
    }

#define TRSV_U_BODY(T, FMA)                                                     \
    for (int i = n - 1; i >= 0; --i) {                                          \
        T s = z[i]; T d = 0; int have = 0;                                      \
        for (int k = rowptr[i] - base; k < rowptr[i + 1] - base; ++k) {         \
            int j = colidx[k] - base;                                           \
            if (j < i) continue;                                                \
            if (j == i) { d = lu[k]; have = 1; continue; }                      \
            s = FMA(-lu[k], y[j], s);                                           \
        }                                                                       \
        if (!have) return ORC_ZERO_PIVOT_STRUCTURAL;                            \
        y[i] = s / d;                                                           \
    }                                                                           \
    return ORC_OK;

void orc_sptrsv_l_d(int n, int base, const int *rowptr, const int *colidx,
                    const double *lu, const double *b, double *z)
{
    TRSV_L_BODY(double, fma)
}
void orc_sptrsv_l_s(int n, int base, const int *rowptr, const int *colidx,
                    const float *lu, const float *b, float *z)
{
    TRSV_L_BODY(float, fmaf)
}
int orc_sptrsv_u_d(int n, int base, const int *rowptr, const int *colidx,
                   const double *lu, const double *z, double *y)
{
    TRSV_U_BODY(double, fma)
}
int orc_sptrsv_u_s(int n, int base, const int *rowptr, const int *colidx,
                   const float *lu, const float *z, float *y)
{
    TRSV_U_BODY(float, fmaf)
}
int orc_sptrsv_lu_d(int n, int base, const int *rowptr, const int *colidx,
                    const double *lu, const double *b, double *z, double *y)
{
    orc_sptrsv_l_d(n, base, rowptr, colidx, lu, b, z);
    return orc_sptrsv_u_d(n, base, rowptr, colidx, lu, z, y);
}
int orc_sptrsv_lu_s(int n, int base, const int *rowptr, const int *colidx,
                    const float *lu, const float *b, float *z, float *y)
{
    orc_sptrsv_l_s(n, base, rowptr, colidx, lu, b, z);
    return orc_sptrsv_u_s(n, base, rowptr, colidx, lu, z, y);
}

/* ---------------------------------------------------------------- levels ---
 * Canonical level sets (what csrilu02_analysis / csrsv2_analysis,
 * ilu0.cu:203-209,232-251, compute internally):
 *   which = 0 (L): lev[i] = 1 + max{lev[j] : j < i, a_ij stored}, 0 if none
 *   which = 1 (U): lev[i] = 1 + max{lev[j] : j > i, a_ij stored}, 0 if none
 * level_ptr has nlevels+1 entries; level_rows lists rows ascending inside a level. */
int orc_levels(int n, int base, const int *rowptr, const int *colidx, int which,
               int *lev, int *nlevels, int *level_ptr, int *level_rows)
{
    int nl = 0;
    if (which == 0) {
        for (int i = 0; i < n; ++i) {
            int l = 0;
            for (int k = rowptr[i] - base; k < rowptr[i + 1] - base; ++k) {
                int j = colidx[k] - base;
                if (j < i && lev[j] + 1 > l) l = lev[j] + 1;
            }
            lev[i] = l;
            if (l + 1 > nl) nl = l + 1;
        }
    } else {
        for (int i = n - 1; i >= 0; --i) {
            int l = 0;
            for (int k = rowptr[i] - base; k < rowptr[i + 1] - base; ++k) {
                int j = colidx[k] - base;
                if (j > i && j < n && lev[j] + 1 > l) l = lev[j] + 1;
            }
            lev[i] = l;
            if (l + 1 > nl) nl = l + 1;
        }
    }
    *nlevels = nl;
    if (level_ptr && level_rows) {
        for (int l = 0; l <= nl; ++l) level_ptr[l] = 0;
        for (int i = 0; i < n; ++i) level_ptr[lev[i] + 1]++;
        for (int l = 0; l < nl; ++l) level_ptr[l + 1] += level_ptr[l];
        int *cur = (int *)malloc(sizeof(int) * (size_t)(nl + 1));
        if (!cur) return ORC_BAD_ARG;
        memcpy(cur, level_ptr, sizeof(int) * (size_t)(nl + 1));
        for (int i = 0; i < n; ++i) level_rows[cur[lev[i]]++] = i;
        free(cur);
    }
    return ORC_OK;
}

/* ------------------------------------------------------------- partition ---
 * nnz-balanced contiguous row blocks for the multi-GPU SpMV:
 * start[g] = first row r with rowptr[r]-base >= floor(g*nnz/P); start[0]=0, start[P]=m.
 * balanced = 0 gives the equal-rows variant ceil(g*m/P).                      */
void orc_row_partition(int m, int base, const int *rowptr, int P, int balanced, int *start)
{
    int64_t nnz = (int64_t)rowptr[m] - rowptr[0];
    start[0] = 0;
    start[P] = m;
    for (int g = 1; g < P; ++g) {
        if (!balanced) {
            start[g] = (int)(((int64_t)g * m + P - 1) / P);
            continue;
        }
        int64_t target = (int64_t)g * nnz / P + (rowptr[0] - base);
        int lo = 0, hi = m;
        while (lo < hi) {
            int mid = lo + (hi - lo) / 2;
            if ((int64_t)rowptr[mid] - base < target) lo = mid + 1; else hi = mid;
        }
        start[g] = lo;
    }
}

/* ------------------------------------------------------------------ PCG ---
 * Textbook preconditioned conjugate gradients (Hestenes-Stiefel; Saad, Iterative Methods, Alg. 9.1)
 * on the oracle's own SpMV and triangular solves -- the checker of rsg_pcg_d (the
 * reference has no Krylov loop, README.md:5-7 names the use).  lu_precision 0: none; 'd': fp64
 * factors; 's': fp32 factors applied to r rounded to fp32, z widened back.  Left-to-right dots.
 * Returns the iteration count; *relres = |r|_2/|b|_2 of the recurrence residual. */
static double orc_dot(int n, const double *a, const double *b)
{
    double s = 0;
    for (int i = 0; i < n; ++i) s = fma(a[i], b[i], s);
    return s;
}

int orc_pcg_d(int n, int base, const int *rowptr, const int *colidx, const double *values,
              const void *lu, int lu_precision, const double *b, double *x, int max_iters, double rtol,
              double *relres)
{
    double *r = malloc(sizeof(double) * (size_t)(n + 1)), *z = malloc(sizeof(double) * (size_t)(n + 1));
    double *p = malloc(sizeof(double) * (size_t)(n + 1)), *q = malloc(sizeof(double) * (size_t)(n + 1));
    float *r32 = malloc(sizeof(float) * (size_t)(n + 1)), *z32 = malloc(sizeof(float) * (size_t)(n + 1));
    float *t32 = malloc(sizeof(float) * (size_t)(n + 1));
    orc_spmv_csr_d(n, base, rowptr, colidx, values, 1.0, x, 0.0, q);
    for (int i = 0; i < n; ++i) r[i] = b[i] - q[i];
    double bb = orc_dot(n, b, b), rr = orc_dot(n, r, r);
    if (!(bb > 0)) bb = 1.0;
    const double stop = rtol * rtol * bb;
    int it = 0;
    double rz = 0;
    if (!(rr <= stop)) {
        for (;;) {
            if (lu_precision == 0) {
                for (int i = 0; i < n; ++i) z[i] = r[i];
            } else if (lu_precision == 's') {
                for (int i = 0; i < n; ++i) r32[i] = (float)r[i];
                orc_sptrsv_lu_s(n, base, rowptr, colidx, (const float *)lu, r32, t32, z32);
                for (int i = 0; i < n; ++i) z[i] = (double)z32[i];
            } else {
                orc_sptrsv_lu_d(n, base, rowptr, colidx, (const double *)lu, r, q, z);
            }
            const double rz_new = orc_dot(n, r, z);
            if (it == 0) {
                for (int i = 0; i < n; ++i) p[i] = z[i];
            } else {
                const double beta = rz_new / rz;
                for (int i = 0; i < n; ++i) p[i] = fma(beta, p[i], z[i]);
            }
            rz = rz_new;
            if (it >= max_iters || !(rr > stop)) break;
            orc_spmv_csr_d(n, base, rowptr, colidx, values, 1.0, p, 0.0, q);
            const double alpha = rz / orc_dot(n, p, q);
            rr = 0;
            for (int i = 0; i < n; ++i) {
                x[i] = fma(alpha, p[i], x[i]);
                r[i] = fma(-alpha, q[i], r[i]);
                rr = fma(r[i], r[i], rr);
            }
            ++it;
        }
    }
    if (relres) *relres = sqrt(rr / bb);
    free(r); free(z); free(p); free(q); free(r32); free(z32); free(t32);
    return it;
}

/* ------------------------------------------------------------- BiCGStab ---
 * van der Vorst's BiCGStab with right preconditioning K = LU (Saad, Alg. 9.7 / Barrett et al.,
 * Templates, fig. 2.10), shadow residual = initial residual; checker of rsg_bicgstab_d.
 * Convergence is tested on |r| once per iteration (no half-step test), like the GPU loop. */
static void orc_apply_prec(int n, int base, const int *rowptr, const int *colidx, const void *lu, int lu_precision,
                           const double *in, double *out, double *tmp, float *in32, float *out32, float *tmp32)
{
    if (lu_precision == 0) {
        for (int i = 0; i < n; ++i) out[i] = in[i];
    } else if (lu_precision == 's') {
        for (int i = 0; i < n; ++i) in32[i] = (float)in[i];
        orc_sptrsv_lu_s(n, base, rowptr, colidx, (const float *)lu, in32, tmp32, out32);
        for (int i = 0; i < n; ++i) out[i] = (double)out32[i];
    } else {
        orc_sptrsv_lu_d(n, base, rowptr, colidx, (const double *)lu, in, tmp, out);
    }
}

int orc_bicgstab_d(int n, int base, const int *rowptr, const int *colidx, const double *values,
                   const void *lu, int lu_precision, const double *b, double *x, int max_iters, double rtol,
                   double *relres)
{
    const size_t nb = sizeof(double) * (size_t)(n + 1);
    double *r = malloc(nb), *rh = malloc(nb), *p = malloc(nb), *v = malloc(nb), *y = malloc(nb), *z = malloc(nb), *t = malloc(nb);
    float *a32 = malloc(nb / 2), *b32 = malloc(nb / 2), *c32 = malloc(nb / 2);
    orc_spmv_csr_d(n, base, rowptr, colidx, values, 1.0, x, 0.0, v);
    for (int i = 0; i < n; ++i) { r[i] = b[i] - v[i]; rh[i] = r[i]; }
    double bb = orc_dot(n, b, b), rr = orc_dot(n, r, r);
    if (!(bb > 0)) bb = 1.0;
    const double stop = rtol * rtol * bb;
    double rho = rr, rho_old = 1, alpha = 1, omega = 1;
    int it = 0;
    while (it < max_iters && rr > stop && isfinite(rr)) {
        if (it == 0) {
            for (int i = 0; i < n; ++i) p[i] = r[i];
        } else {
            const double beta = (rho / rho_old) * (alpha / omega);
            for (int i = 0; i < n; ++i) p[i] = fma(beta, fma(-omega, v[i], p[i]), r[i]);
        }
        orc_apply_prec(n, base, rowptr, colidx, lu, lu_precision, p, y, t, a32, b32, c32);
        orc_spmv_csr_d(n, base, rowptr, colidx, values, 1.0, y, 0.0, v);
        alpha = rho / orc_dot(n, rh, v);
        for (int i = 0; i < n; ++i) r[i] = fma(-alpha, v[i], r[i]); /* s */
        orc_apply_prec(n, base, rowptr, colidx, lu, lu_precision, r, z, t, a32, b32, c32);
        orc_spmv_csr_d(n, base, rowptr, colidx, values, 1.0, z, 0.0, t);
        omega = orc_dot(n, t, r) / orc_dot(n, t, t);
        rr = 0;
        double rho_next = 0;
        for (int i = 0; i < n; ++i) {
            x[i] = fma(omega, z[i], fma(alpha, y[i], x[i]));
            r[i] = fma(-omega, t[i], r[i]);
            rr = fma(r[i], r[i], rr);
            rho_next = fma(rh[i], r[i], rho_next);
        }
        rho_old = rho;
        rho = rho_next;
        ++it;
    }
    if (relres) *relres = sqrt(rr / bb);
    free(r); free(rh); free(p); free(v); free(y); free(z); free(t); free(a32); free(b32); free(c32);
    return it;
}

/* ---------------------------------------------------------------- GMRES ---
 * Restarted flexible GMRES(30) with right preconditioning K = LU (Saad & Schultz 1986; Saad, Iterative Methods,
 * Alg. 9.6: the preconditioned basis Z is kept, so an fp32 preconditioner -- not a linear operator once its input is
 * rounded -- costs no accuracy), modified Gram-Schmidt, Givens rotations on the host-size Hessenberg matrix; checker of rsg_gmres_d.  One
 * iteration = one Arnoldi step; |r| is the rotated right-hand side's last entry, the true residual is recomputed
 * at every restart.  The reference has no Krylov loop (README.md:5-7 names the use only). */
#define ORC_GMRES_RESTART 30
int orc_gmres_d(int n, int base, const int *rowptr, const int *colidx, const double *values,
                const void *lu, int lu_precision, const double *b, double *x, int max_iters, double rtol,
                double *relres)
{
    enum { R = ORC_GMRES_RESTART };
    const size_t nd = (size_t)n + 1;
    double *V = malloc(sizeof(double) * nd * (R + 1)), *w = malloc(sizeof(double) * nd), *Z = malloc(sizeof(double) * nd * R),
           *t = malloc(sizeof(double) * nd);
    float *a32 = malloc(sizeof(float) * nd), *b32 = malloc(sizeof(float) * nd), *c32 = malloc(sizeof(float) * nd);
    double H[R + 1][R], cs[R], sn[R], g[R + 1], yv[R];
    double bb = orc_dot(n, b, b), rr = 0;
    if (!(bb > 0)) bb = 1.0;
    const double stop = rtol * rtol * bb;
    int it = 0;
    for (;;) {
        orc_spmv_csr_d(n, base, rowptr, colidx, values, 1.0, x, 0.0, w);
        for (int i = 0; i < n; ++i) V[i] = b[i] - w[i];
        rr = orc_dot(n, V, V);
        if (!(rr > stop) || !isfinite(rr) || it >= max_iters) break;
        const double inv0 = 1.0 / sqrt(rr);
        for (int i = 0; i < n; ++i) V[i] *= inv0;
        g[0] = sqrt(rr);
        for (int i = 1; i <= R; ++i) g[i] = 0;
        int k = 0;
        for (int j = 0; j < R && it < max_iters; ++j) {
            double *vj = V + nd * j;
            double *z = Z + nd * j; /* flexible variant: z_j = K^-1 v_j is kept, x += Z y */
            orc_apply_prec(n, base, rowptr, colidx, lu, lu_precision, vj, z, t, a32, b32, c32);
            orc_spmv_csr_d(n, base, rowptr, colidx, values, 1.0, z, 0.0, w);
            for (int i = 0; i <= j; ++i) {
                const double *vi = V + nd * i;
                const double h = orc_dot(n, w, vi);
                H[i][j] = h;
                for (int q = 0; q < n; ++q) w[q] = fma(-h, vi[q], w[q]);
            }
            const double ww = orc_dot(n, w, w), hn = sqrt(ww);
            for (int i = 0; i < j; ++i) {
                const double tmp = cs[i] * H[i][j] + sn[i] * H[i + 1][j];
                H[i + 1][j] = -sn[i] * H[i][j] + cs[i] * H[i + 1][j];
                H[i][j] = tmp;
            }
            const double d = hypot(H[j][j], hn);
            if (d == 0) break;
            cs[j] = H[j][j] / d;
            sn[j] = hn / d;
            H[j][j] = d;
            g[j + 1] = -sn[j] * g[j];
            g[j] = cs[j] * g[j];
            ++it;
            k = j + 1;
            rr = g[j + 1] * g[j + 1];
            if (!(rr > stop) || !isfinite(rr) || hn == 0) break;
            if (j + 1 < R && it < max_iters) {
                const double inv = 1.0 / sqrt(ww);
                double *vn = V + nd * (j + 1);
                for (int q = 0; q < n; ++q) vn[q] = w[q] * inv;
            }
        }
        for (int i = k - 1; i >= 0; --i) {
            double s = g[i];
            for (int q = i + 1; q < k; ++q) s -= H[i][q] * yv[q];
            yv[i] = s / H[i][i];
        }
        for (int q = 0; q < n; ++q) {
            double s = 0;
            for (int i = 0; i < k; ++i) s = fma(yv[i], Z[nd * i + q], s);
            x[q] += s;
        }
        if (!(rr > stop) || !isfinite(rr) || it >= max_iters || k == 0) break;
    }
    if (relres) *relres = sqrt(rr / bb);
    free(V); free(w); free(Z); free(t); free(a32); free(b32); free(c32);
    return it;
}
