// test_spmv -- drop-in for the reference's GPU SpMV harness (/root/reference/GPU/spmv.cu, binary
// `test_spmv`, GPU/Makefile:10): same positional argument, same two stdout lines
//   "DOUBLE|SINGLE PRECISION SPMV solve time (microseconds) = %f"   and   "Error= %e",
// with the cusparseSpMV call group replaced by libreaspasol_gpu's spmv_csr_d / _s / _mixed and
// the MKL cross-check replaced by a deterministic sequential host product.
// Extra flags (all optional): --precision=d|s|m --repeat=N --warmup=N --kernel=auto|window|tile|rowblock|merge|vector|scalar --stats
//   --gpus=N   the same product row-block partitioned over N devices of this process (rsg_spmv_multi_*): same two lines
#include <cmath>

#include "harness_common.h"

template <typename TV, typename TX>
static int run(const HarnessArgs &args, const CSR &A, int (*spmv)(rsg_handle_t, int, int, int, int, const int *, const int *,
                                                                  const TV *, TX, const TX *, TX, TX *))
{
    // The reference assumes a square matrix and uses A.n for both (spmv.cu:50-51); keep m and n apart.
    const int m = A.m, n = A.n, stored = rsg_csr_stored_nnz(&A);
    std::vector<TV> vals((size_t)stored);
#pragma omp parallel for
    for (int i = 0; i < stored; ++i) vals[(size_t)i] = (TV)A.values[i]; // fp32 cast as in spmv.cu:63-67
    std::vector<TX> X((size_t)n, (TX)1), Y((size_t)m, (TX)0);          // x = ones (spmv.cu:71,83)

    int *dRow = to_device(A.rowptr, (size_t)m + 1), *dCol = to_device(A.colidx, (size_t)stored);
    TV *dVal = to_device(vals.data(), (size_t)stored);
    TX *dX = to_device(X.data(), (size_t)n), *dY = to_device(Y.data(), (size_t)m);

    rsg_handle_t h = nullptr;
    HARNESS_RSG(rsg_create(&h, 0));
    HARNESS_RSG(rsg_set_option(h, RSG_OPT_SPMV_ALGO, args.algo));
    // analysis = the analogue of cusparseSpMV_bufferSize + workspace (spmv.cu:159-164): untimed
    HARNESS_RSG(rsg_spmv_analysis(h, m, n, stored, 0, dRow, dCol));

    cudaEvent_t start, stop;
    HARNESS_CUDA(cudaEventCreate(&start));
    HARNESS_CUDA(cudaEventCreate(&stop));
    for (int i = 0; i < args.warmup; ++i) HARNESS_RSG(spmv(h, m, n, stored, 0, dRow, dCol, dVal, (TX)1, dX, (TX)0, dY));
    std::vector<float> times((size_t)args.repeat);
    for (int i = 0; i < args.repeat; ++i) {
        HARNESS_CUDA(cudaEventRecord(start));
        HARNESS_RSG(spmv(h, m, n, stored, 0, dRow, dCol, dVal, (TX)1, dX, (TX)0, dY));
        HARNESS_CUDA(cudaEventRecord(stop));
        HARNESS_CUDA(cudaEventSynchronize(stop));
        HARNESS_CUDA(cudaEventElapsedTime(&times[(size_t)i], start, stop));
    }
    float sum = 0.0f, best = times[0];
    for (float t : times) { sum += t; best = t < best ? t : best; }
    printf(sizeof(TX) == 4 ? "SINGLE PRECISION SPMV " : "DOUBLE PRECISION SPMV ");
    printf("solve time (microseconds) = %f\n", (sum / args.repeat) * 1000);

    HARNESS_CUDA(cudaMemcpy(Y.data(), dY, sizeof(TX) * (size_t)m, cudaMemcpyDeviceToHost));
    // cross-check (the reference uses mkl_sparse_?_mv here, spmv.cu:221-254) with a proper reduction
    double error = 0.0;
#pragma omp parallel for reduction(+ : error)
    for (int i = 0; i < m; ++i) {
        TX s = 0;
        for (int k = A.rowptr[i]; k < A.rowptr[i + 1]; ++k) s += (TX)vals[(size_t)k] * X[(size_t)A.colidx[k]];
        error += std::fabs((double)s - (double)Y[(size_t)i]);
    }
    printf("Error= %e\n", error / n);
    if (args.stats) {
        const double bytes = (double)stored * (sizeof(TV) + 4) + 4.0 * (m + 1) + (double)sizeof(TX) * ((double)m + n);
        long long algo = 0, blocks = 0;
        rsg_get_info(h, RSG_INFO_SPMV_ALGO_USED, &algo);
        rsg_get_info(h, RSG_INFO_SPMV_ROW_BLOCKS, &blocks);
        printf("rsg: m=%d n=%d stored=%d (reported nnz=%d) algo=%lld ctas=%lld best=%.3f us mean=%.3f us  %.1f GB/s  %.1f GFLOP/s (best)\n",
               m, n, stored, A.nnz, algo, blocks, best * 1000, sum / args.repeat * 1000, bytes / (best * 1e-3) * 1e-9,
               2.0 * stored / (best * 1e-3) * 1e-9);
    }
    rsg_destroy(h);
    cudaFree(dRow); cudaFree(dCol); cudaFree(dVal); cudaFree(dX); cudaFree(dY);
    cudaEventDestroy(start); cudaEventDestroy(stop);
    return 0;
}

// --gpus=N: the whole matrix over N devices (row blocks, x halo pushed over NVLink inside each product's launch).
// The timed loop keeps the reference's shape (spmv.cu:174-207: REPEAT products, mean time), measured on the devices.
template <typename TX> static int run_multi(const HarnessArgs &args, const CSR &A)
{
    const int m = A.m, n = A.n, stored = rsg_csr_stored_nnz(&A);
    if (m != n) { fprintf(stderr, "--gpus needs a square matrix (x is distributed like the rows)\n"); return -1; }
    std::vector<TX> X((size_t)n, (TX)1), Y((size_t)m, (TX)0);
    rsg_multi_t M = nullptr;
    int rc = rsg_spmv_multi_create(args.gpus, nullptr, args.precision, m, n, stored, 0, A.rowptr, A.colidx, A.values, nullptr, &M);
    if (rc != RSG_OK) { fprintf(stderr, "rsg_spmv_multi_create: %s\n", rsg_status_string(rc)); return -1; }
    HARNESS_RSG(rsg_spmv_multi_set_x(M, X.data()));
    if (args.warmup > 0) HARNESS_RSG(rsg_spmv_multi_step(M, args.warmup));
    double us = 0.0;
    HARNESS_RSG(rsg_spmv_multi_time(M, args.repeat, &us));
    printf(sizeof(TX) == 4 ? "SINGLE PRECISION SPMV " : "DOUBLE PRECISION SPMV ");
    printf("solve time (microseconds) = %f\n", us);
    HARNESS_RSG(rsg_spmv_multi_get_y(M, Y.data()));
    double error = 0.0;
    const bool f32 = args.precision != 'd';
#pragma omp parallel for reduction(+ : error)
    for (int i = 0; i < m; ++i) {
        TX s = 0;
        for (int k = A.rowptr[i]; k < A.rowptr[i + 1]; ++k) s += (TX)(f32 ? (double)(float)A.values[k] : A.values[k]) * X[(size_t)A.colidx[k]];
        error += std::fabs((double)s - (double)Y[(size_t)i]);
    }
    printf("Error= %e\n", error / n);
    if (args.stats) {
        long long halo = 0, algo = 0;
        std::vector<int> cuts((size_t)args.gpus + 1);
        rsg_spmv_multi_info(M, 1, &halo, cuts.data());
        rsg_spmv_multi_info(M, 2, &algo, nullptr);
        const double vb = args.precision == 'd' ? 8 : 4, xb = sizeof(TX);
        const double bytes = (double)stored * (vb + 4) + 4.0 * (m + 1) + xb * ((double)m + n);
        printf("rsg: gpus=%d m=%d stored=%d algo=%lld halo_bytes_per_step=%lld mean=%.3f us  %.1f GB/s aggregate  cuts:", args.gpus, m, stored,
               algo, halo, us, bytes / (us * 1e-6) * 1e-9);
        for (int c : cuts) printf(" %d", c);
        printf("\n");
    }
    rsg_spmv_multi_destroy(M);
    return 0;
}

int main(int argc, char **argv)
{
    HarnessArgs args;
    if (!parse_args(argc, argv, &args)) return -1;
    CSR A;
    if (!rsg_load_matrix_market(args.matrix, &A, 0, 0, args.expand_symmetric ? RSG_LOADER_EXPAND_SYMMETRIC : RSG_LOADER_BUGCOMPAT))
        return -1;
    int rc;
    if (args.gpus > 1) rc = args.precision == 's' ? run_multi<float>(args, A) : run_multi<double>(args, A);
    else if (args.precision == 's') rc = run<float, float>(args, A, spmv_csr_s);
    else if (args.precision == 'm') rc = run<float, double>(args, A, spmv_csr_mixed);
    else rc = run<double, double>(args, A, spmv_csr_d);
    rsg_free_csr(&A);
    return rc;
}
