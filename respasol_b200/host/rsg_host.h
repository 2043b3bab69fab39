/*
 * rsg_host.h -- host-side companions of libreaspasol_gpu: the Matrix-Market -> CSR loader
 * (drop-in for the reference's ReadMatrixMarket library) and the synthetic matrix
 * generators that stand in for the SuiteSparse downloads (no network).
 *
 * Reference interfaces mirrored (paths relative to /root/reference):
 *   struct CSR                     ReadMatrixMarket/loadMatrixMarket.h:17-25 (same field order/types)
 *   int loadMatrixMarket(...)      ReadMatrixMarket/loadMatrixMarket.h:41, .cpp:47-253
 *   MALLOC/FREE 64-byte alignment  ReadMatrixMarket/loadMatrixMarket.h:14-15
 * Built as librsg_host.so (respasol_b200/csrc/Makefile).  No CUDA in here.
 */
#ifndef RSG_HOST_H
#define RSG_HOST_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Layout-identical to the reference CSR so harness code that fills or reads one is unchanged. */
typedef struct {
    int isSymmetric;
    int m;
    int n;
    int nnz;
    int *rowptr;
    int *colidx;
    double *values;
} CSR;

/* loader flags */
#define RSG_LOADER_BUGCOMPAT 0        /* default: bit-identical to the reference, including its
                                         dropped mirrored half of `symmetric` files */
#define RSG_LOADER_EXPAND_SYMMETRIC 1 /* fixed mode: mirrored entries are really inserted and
                                         nnz == rowptr[m]-rowptr[0] */

/* Returns 1 on success, 0 on failure (the reference's convention); never exit()s.
 * outputBase 0 => zero-based indices (follows the code, not the swapped docstring, 9-D6);
 * transpose 1 => CSC of the file's matrix (loadMatrixMarket.cpp:79-81,124). */
int rsg_load_matrix_market(const char *file, CSR *matrix, int outputBase, int transpose, int flags);

/* Effective number of stored entries, rowptr[m]-rowptr[0]; differs from matrix->nnz on
 * symmetric files loaded in bug-compatible mode. */
int rsg_csr_stored_nnz(const CSR *matrix);

void rsg_free_csr(CSR *matrix);

/* Writes `coordinate real general`, 1-based, "%d %d %.17g".  base = index base of the CSR. */
int rsg_write_matrix_market(const char *file, const CSR *matrix, int base);

/* ---- synthetic generators; all emit base-0 CSR, sorted unique columns ---- */

/* C1: 5-point Laplacian on an N x N grid, diag 4, off-diag -1; nnz = 5N^2-4N. */
int rsg_gen_laplace2d(int N, CSR *out);

/* C2/C5: 27-point stencil on nx*ny*nz, rows of planes [z0,z1) only (global column indices,
 * out->n = nx*ny*nz).  diag 26, off-diag -1*(1+0.01u), u~U(-1,1) hashed from (seed,row,col);
 * seed 0 => exactly -1.  Full grid: z0=0,z1=nz => nnz=(3nx-2)(3ny-2)(3nz-2). */
int rsg_gen_stencil27(int nx, int ny, int nz, int z0, int z1, uint64_t seed, CSR *out);

/* C3: FEM-like for ILU0: N^3 nodes, 27-point connectivity, each off-diagonal PAIR kept with
 * probability p (hash of seed,lo,hi), values -(0.5+u), diagonal 1.05*sum|offdiag| (strictly
 * diagonally dominant => ILU0 exists without pivoting). */
int rsg_gen_femlike(int N, double p, uint64_t seed, CSR *out);

/* C4: power-law rows: m rows, length 1+floor(Pareto(alpha)) clipped to [1,maxlen] random distinct
 * columns + diagonal, plus ndense dense rows (and their transposed columns) of dense_len entries. */
int rsg_gen_powerlaw(int m, double alpha, int maxlen, int ndense, int dense_len, uint64_t seed, CSR *out);

/* Generic shape class (sweep harness): m rows x per_row entries (diagonal + hashed distinct columns within
 * +-half_band), diagonally dominant values. */
int rsg_gen_banded(int m, int per_row, int half_band, uint64_t seed, CSR *out);

/* dlarnv(1,{0,0,0,1}) restated for the harness (test_spmv.c:75-76): 48-bit LCG, u = s*2^-48. */
void rsg_dlarnv_u01(int64_t n, double *x);

#ifdef __cplusplus
}

/* Reference-signature wrapper (C++ linkage, like loadMatrixMarket.h:41) for drop-in harness code. */
static inline int loadMatrixMarket(const char *file, CSR *matrix, int outputBase, int transpose)
{
    return rsg_load_matrix_market(file, matrix, outputBase, transpose, RSG_LOADER_BUGCOMPAT);
}
#endif

#endif /* RSG_HOST_H */
