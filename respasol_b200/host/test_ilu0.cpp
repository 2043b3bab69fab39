// test_ilu0 -- drop-in for the reference's GPU ILU0 harness (/root/reference/GPU/ilu0.cu, binary
// `test_ilu0`, GPU/Makefile:14): same positional argument, same stdout block
//   "DOUBLE|SINGLE PRECISION SOLVE IN  MILLISECONDS\n Symbolic = ..\n Numeric = .. \n Symbolic+ Numeric = ..\n Solve = ..",
// same zero-pivot messages and exit codes (ilu0.cu:221-226,278-282).  The csrilu02 / csrsv2 call
// groups are replaced by rsg_ilu0_analysis, ilu0_factor_*, rsg_sptrsv_analysis (untimed, as in the
// reference) and sptrsv_lu_*; the second solve is the intended U solve.
// Unlike the reference (no result check at all) --stats prints the residual of the solves.
#include <cmath>

#include "harness_common.h"

template <typename T>
static int run(const HarnessArgs &args, const CSR &A,
               int (*factor)(rsg_handle_t, int, int, int, const int *, const int *, T *, int *),
               int (*solve)(rsg_handle_t, int, int, int, const int *, const int *, const T *, const T *, T *, T *))
{
    const int n = A.n, stored = rsg_csr_stored_nnz(&A);
    if (A.m != A.n) { fprintf(stderr, "ILU0 needs a square matrix (%d x %d)\n", A.m, A.n); return -1; }
    std::vector<T> vals((size_t)stored), X((size_t)n, (T)1), Y((size_t)n, (T)0);
#pragma omp parallel for
    for (int i = 0; i < stored; ++i) vals[(size_t)i] = (T)A.values[i];
    int *dRow = to_device(A.rowptr, (size_t)n + 1), *dCol = to_device(A.colidx, (size_t)stored);
    T *dVal = to_device(vals.data(), (size_t)stored);
    T *dX = to_device(X.data(), (size_t)n), *dY = to_device(Y.data(), (size_t)n), *dZ = to_device(Y.data(), (size_t)n);

    rsg_handle_t h = nullptr;
    HARNESS_RSG(rsg_create(&h, 0));
    cudaEvent_t start, stop;
    HARNESS_CUDA(cudaEventCreate(&start));
    HARNESS_CUDA(cudaEventCreate(&stop));
    float t_sym = 0, t_num = 0, t_solve = 0;

    HARNESS_CUDA(cudaEventRecord(start));
    int pivot = -1;
    int st = rsg_ilu0_analysis(h, n, stored, 0, dRow, dCol, &pivot);
    HARNESS_CUDA(cudaEventRecord(stop));
    HARNESS_CUDA(cudaEventSynchronize(stop));
    HARNESS_CUDA(cudaEventElapsedTime(&t_sym, start, stop));
    if (st == RSG_ZERO_PIVOT_STRUCTURAL) {
        printf("A(%d,%d) is missing\n", pivot, pivot);
        return 0;
    }
    HARNESS_RSG(st);
    HARNESS_RSG(rsg_sptrsv_analysis(h, n, stored, 0, dRow, dCol)); // untimed in the reference (ilu0.cu:229-252)

    HARNESS_CUDA(cudaEventRecord(start));
    st = factor(h, n, stored, 0, dRow, dCol, dVal, &pivot);
    HARNESS_CUDA(cudaEventRecord(stop));
    HARNESS_CUDA(cudaEventSynchronize(stop));
    HARNESS_CUDA(cudaEventElapsedTime(&t_num, start, stop));
    if (st == RSG_ZERO_PIVOT_NUMERIC) {
        printf("L(%d,%d) is zero\n", pivot, pivot);
        return 0;
    }
    HARNESS_RSG(st);

    HARNESS_CUDA(cudaEventRecord(start));
    HARNESS_RSG(solve(h, n, stored, 0, dRow, dCol, dVal, dX, dZ, dY)); // L z = x, then U y = z
    HARNESS_CUDA(cudaEventRecord(stop));
    HARNESS_CUDA(cudaEventSynchronize(stop));
    HARNESS_CUDA(cudaEventElapsedTime(&t_solve, start, stop));

    printf(sizeof(T) == 4 ? "SINGLE PRECISION SOLVE IN  MILLISECONDS\n " : "DOUBLE PRECISION SOLVE IN  MILLISECONDS\n ");
    printf("Symbolic = %f\n Numeric = %f \n Symbolic+ Numeric = %f\n Solve = %f\n", t_sym, t_num, t_sym + t_num, t_solve);

    if (args.stats) {
        std::vector<T> lu((size_t)stored);
        HARNESS_CUDA(cudaMemcpy(lu.data(), dVal, sizeof(T) * (size_t)stored, cudaMemcpyDeviceToHost));
        HARNESS_CUDA(cudaMemcpy(Y.data(), dY, sizeof(T) * (size_t)n, cudaMemcpyDeviceToHost));
        // residual of the two solves in fp64: r = L (U y) - x
        std::vector<double> w((size_t)n), r((size_t)n);
        for (int i = 0; i < n; ++i) {
            double s = 0;
            for (int k = A.rowptr[i]; k < A.rowptr[i + 1]; ++k)
                if (A.colidx[k] >= i) s += (double)lu[(size_t)k] * (double)Y[(size_t)A.colidx[k]];
            w[(size_t)i] = s;
        }
        double rmax = 0;
        for (int i = 0; i < n; ++i) {
            double s = w[(size_t)i];
            for (int k = A.rowptr[i]; k < A.rowptr[i + 1]; ++k)
                if (A.colidx[k] < i) s += (double)lu[(size_t)k] * w[(size_t)A.colidx[k]];
            rmax = std::fmax(rmax, std::fabs(s - 1.0));
        }
        long long ll = 0, lu_levels = 0;
        rsg_get_info(h, RSG_INFO_LEVELS_L, &ll);
        rsg_get_info(h, RSG_INFO_LEVELS_U, &lu_levels);
        printf("rsg: n=%d stored=%d levels(L)=%lld levels(U)=%lld  ||L U y - x||inf = %e\n", n, stored, ll, lu_levels, rmax);
    }
    rsg_destroy(h);
    cudaFree(dRow); cudaFree(dCol); cudaFree(dVal); cudaFree(dX); cudaFree(dY); cudaFree(dZ);
    cudaEventDestroy(start); cudaEventDestroy(stop);
    return 0;
}

int main(int argc, char **argv)
{
    HarnessArgs args;
    if (!parse_args(argc, argv, &args)) return -1;
    CSR A;
    if (!rsg_load_matrix_market(args.matrix, &A, 0, 0, args.expand_symmetric ? RSG_LOADER_EXPAND_SYMMETRIC : RSG_LOADER_BUGCOMPAT))
        return -1;
    int rc = args.precision == 's' ? run<float>(args, A, ilu0_factor_s, sptrsv_lu_s) : run<double>(args, A, ilu0_factor_d, sptrsv_lu_d);
    rsg_free_csr(&A);
    return rc;
}
