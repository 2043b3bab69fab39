// cusparse_ref.cu -- CHECKER / COMPETITOR, never the product: the reference harness's cuSPARSE call
// groups restated for the cuSPARSE 12.x that is on this box, behind a tiny C ABI so the GPU tests and
// bench.py can (1) compare libreaspasol_gpu's results with the library the reference itself computes
// with, on the same device arrays, and (2) time that library next to ours on the same B200.
//
//   cref_spmv      cusparseCreateCsr/CreateDnVec/SpMV_bufferSize/SpMV         GPU/spmv.cu:132-186
//                  (CUSPARSE_MV_ALG_DEFAULT was renamed CUSPARSE_SPMV_ALG_DEFAULT; alg selectable;
//                   cusparseSpMV_preprocess is run once, untimed, like the buffer query)
//   cref_ilu0      csrilu02_bufferSize/_analysis/zeroPivot/csrilu02/zeroPivot GPU/ilu0.cu:166-282
//   cref_spsv_lu   csrsv2 was removed in CUDA 12: cusparseSpSV_analysis/solve, L = lower/unit,
//                  U = upper/non-unit on the same LU values (the INTENDED second solve)
//                                                                             GPU/ilu0.cu:229-310
// All pointers are DEVICE pointers, int32 base-0 indices.  Timings use cudaEvents on the default
// stream exactly like the harness (spmv.cu:176-195, ilu0.cu:199-310).
// Only tests/, smoke() and bench.py may load this library (oracle/ rule).
#include <cuda_runtime.h>
#include <cusparse.h>
#include <algorithm>
#include <vector>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) return 1000 + (int)e_; } while (0)
#define SP(x) do { cusparseStatus_t s_ = (x); if (s_ != CUSPARSE_STATUS_SUCCESS) return 2000 + (int)s_; } while (0)

namespace {
struct Timer {
    cudaEvent_t a, b;
    Timer() { cudaEventCreate(&a); cudaEventCreate(&b); }
    ~Timer() { cudaEventDestroy(a); cudaEventDestroy(b); }
    void start() { cudaEventRecord(a, 0); }
    float stop() { cudaEventRecord(b, 0); cudaEventSynchronize(b); float ms = 0; cudaEventElapsedTime(&ms, a, b); return ms; }
};
}  // namespace

extern "C" int cref_version(void) { return CUSPARSE_VERSION; }

// prec: 0 = fp64, 1 = fp32.  alg: 0 default, 1 CSR_ALG1, 2 CSR_ALG2.  flush/flush_bytes: optional device
// scratch written between repetitions (outside the event pair) so every product starts with a cold L2.
// us[0] = mean, us[1] = median, us[2] = min over `reps` products (after `warm` untimed ones).
static int spmv_any(int transpose, int prec, int alg, int m, int n, int nnz, const int *rowptr, const int *colidx,
                    const void *values, const void *x, void *y, int warm, int reps, void *flush,
                    size_t flush_bytes, float *us);

extern "C" int cref_spmv(int prec, int alg, int m, int n, int nnz, const int *rowptr, const int *colidx,
                         const void *values, const void *x, void *y, int warm, int reps, void *flush,
                         size_t flush_bytes, float *us)
{
    return spmv_any(0, prec, alg, m, n, nnz, rowptr, colidx, values, x, y, warm, reps, flush, flush_bytes, us);
}

// y (n entries) = A^T x (m entries): CUSPARSE_OPERATION_TRANSPOSE of the same call group
extern "C" int cref_spmv_t(int prec, int alg, int m, int n, int nnz, const int *rowptr, const int *colidx,
                           const void *values, const void *x, void *y, int warm, int reps, void *flush,
                           size_t flush_bytes, float *us)
{
    return spmv_any(1, prec, alg, m, n, nnz, rowptr, colidx, values, x, y, warm, reps, flush, flush_bytes, us);
}

static int spmv_any(int transpose, int prec, int alg, int m, int n, int nnz, const int *rowptr, const int *colidx,
                    const void *values, const void *x, void *y, int warm, int reps, void *flush,
                    size_t flush_bytes, float *us)
{
    const cusparseOperation_t op = transpose ? CUSPARSE_OPERATION_TRANSPOSE : CUSPARSE_OPERATION_NON_TRANSPOSE;
    cusparseHandle_t h;
    SP(cusparseCreate(&h));
    cudaDataType dt = prec ? CUDA_R_32F : CUDA_R_64F;
    cusparseSpMatDescr_t A;
    cusparseDnVecDescr_t vx, vy;
    SP(cusparseCreateCsr(&A, m, n, nnz, (void *)rowptr, (void *)colidx, (void *)values, CUSPARSE_INDEX_32I,
                         CUSPARSE_INDEX_32I, CUSPARSE_INDEX_BASE_ZERO, dt));
    SP(cusparseCreateDnVec(&vx, transpose ? m : n, (void *)x, dt));
    SP(cusparseCreateDnVec(&vy, transpose ? n : m, y, dt));
    double ad = 1.0, bd = 0.0;
    float af = 1.0f, bf = 0.0f;
    const void *alpha = prec ? (const void *)&af : (const void *)&ad;
    const void *beta = prec ? (const void *)&bf : (const void *)&bd;
    cusparseSpMVAlg_t a = alg == 1 ? CUSPARSE_SPMV_CSR_ALG1 : alg == 2 ? CUSPARSE_SPMV_CSR_ALG2 : CUSPARSE_SPMV_ALG_DEFAULT;
    size_t bs = 0;
    SP(cusparseSpMV_bufferSize(h, op, alpha, A, vx, beta, vy, dt, a, &bs));
    void *buf = nullptr;
    CK(cudaMalloc(&buf, bs ? bs : 16));
    SP(cusparseSpMV_preprocess(h, op, alpha, A, vx, beta, vy, dt, a, buf));
    if (warm < 1 && reps < 1) warm = 1; // a plain product when no timing is asked for
    for (int i = 0; i < warm; ++i)
        SP(cusparseSpMV(h, op, alpha, A, vx, beta, vy, dt, a, buf));
    CK(cudaDeviceSynchronize());
    std::vector<float> t(reps > 0 ? reps : 0);
    Timer tm;
    for (int i = 0; i < reps; ++i) {
        if (flush) CK(cudaMemsetAsync(flush, i & 0xff, flush_bytes, 0));
        tm.start();
        SP(cusparseSpMV(h, op, alpha, A, vx, beta, vy, dt, a, buf));
        t[i] = tm.stop() * 1000.0f;
    }
    if (reps > 0 && us) {
        double s = 0;
        for (float v : t) s += v;
        std::sort(t.begin(), t.end());
        us[0] = (float)(s / reps);
        us[1] = t[reps / 2];
        us[2] = t[0];
    }
    cudaFree(buf);
    cusparseDestroyDnVec(vx);
    cusparseDestroyDnVec(vy);
    cusparseDestroySpMat(A);
    cusparseDestroy(h);
    return 0;
}

// In-place ILU(0) through csrilu02 (policy as ilu0.cu:186 NO_LEVEL when level == 0, USE_LEVEL otherwise).
// pivots[0] = structural zero-pivot row or -1, pivots[1] = numerical.  ms[0] = analysis, ms[1] = numeric.
extern "C" int cref_ilu0(int prec, int level, int n, int nnz, const int *rowptr, const int *colidx, void *values,
                         int *pivots, float *ms)
{
    cusparseHandle_t h;
    SP(cusparseCreate(&h));
    cusparseMatDescr_t d;
    SP(cusparseCreateMatDescr(&d));
    cusparseSetMatIndexBase(d, CUSPARSE_INDEX_BASE_ZERO);
    cusparseSetMatType(d, CUSPARSE_MATRIX_TYPE_GENERAL);
    csrilu02Info_t info;
    SP(cusparseCreateCsrilu02Info(&info));
    cusparseSolvePolicy_t pol = level ? CUSPARSE_SOLVE_POLICY_USE_LEVEL : CUSPARSE_SOLVE_POLICY_NO_LEVEL;
    int bs = 0;
    if (prec) SP(cusparseScsrilu02_bufferSize(h, n, nnz, d, (float *)values, rowptr, colidx, info, &bs));
    else SP(cusparseDcsrilu02_bufferSize(h, n, nnz, d, (double *)values, rowptr, colidx, info, &bs));
    void *buf = nullptr;
    CK(cudaMalloc(&buf, bs > 0 ? bs : 16));
    Timer tm;
    tm.start();
    if (prec) SP(cusparseScsrilu02_analysis(h, n, nnz, d, (float *)values, rowptr, colidx, info, pol, buf));
    else SP(cusparseDcsrilu02_analysis(h, n, nnz, d, (double *)values, rowptr, colidx, info, pol, buf));
    float t0 = tm.stop();
    int piv = -1;
    pivots[0] = pivots[1] = -1;
    if (cusparseXcsrilu02_zeroPivot(h, info, &piv) == CUSPARSE_STATUS_ZERO_PIVOT) pivots[0] = piv;
    tm.start();
    if (prec) SP(cusparseScsrilu02(h, n, nnz, d, (float *)values, rowptr, colidx, info, pol, buf));
    else SP(cusparseDcsrilu02(h, n, nnz, d, (double *)values, rowptr, colidx, info, pol, buf));
    float t1 = tm.stop();
    if (cusparseXcsrilu02_zeroPivot(h, info, &piv) == CUSPARSE_STATUS_ZERO_PIVOT) pivots[1] = piv;
    if (ms) { ms[0] = t0; ms[1] = t1; }
    cudaFree(buf);
    cusparseDestroyCsrilu02Info(info);
    cusparseDestroyMatDescr(d);
    cusparseDestroy(h);
    return 0;
}

// L z = b (lower, unit diagonal), U y = z (upper, non-unit) on in-place LU values.
// ms[0] = both analyses, ms[1] = both solves (median of `reps`).
extern "C" int cref_spsv_lu(int prec, int n, int nnz, const int *rowptr, const int *colidx, const void *lu,
                            const void *b, void *z, void *y, int reps, float *ms)
{
    cusparseHandle_t h;
    SP(cusparseCreate(&h));
    cudaDataType dt = prec ? CUDA_R_32F : CUDA_R_64F;
    cusparseSpMatDescr_t L, U;
    SP(cusparseCreateCsr(&L, n, n, nnz, (void *)rowptr, (void *)colidx, (void *)lu, CUSPARSE_INDEX_32I,
                         CUSPARSE_INDEX_32I, CUSPARSE_INDEX_BASE_ZERO, dt));
    SP(cusparseCreateCsr(&U, n, n, nnz, (void *)rowptr, (void *)colidx, (void *)lu, CUSPARSE_INDEX_32I,
                         CUSPARSE_INDEX_32I, CUSPARSE_INDEX_BASE_ZERO, dt));
    cusparseFillMode_t lo = CUSPARSE_FILL_MODE_LOWER, up = CUSPARSE_FILL_MODE_UPPER;
    cusparseDiagType_t un = CUSPARSE_DIAG_TYPE_UNIT, nu = CUSPARSE_DIAG_TYPE_NON_UNIT;
    SP(cusparseSpMatSetAttribute(L, CUSPARSE_SPMAT_FILL_MODE, &lo, sizeof lo));
    SP(cusparseSpMatSetAttribute(L, CUSPARSE_SPMAT_DIAG_TYPE, &un, sizeof un));
    SP(cusparseSpMatSetAttribute(U, CUSPARSE_SPMAT_FILL_MODE, &up, sizeof up));
    SP(cusparseSpMatSetAttribute(U, CUSPARSE_SPMAT_DIAG_TYPE, &nu, sizeof nu));
    cusparseDnVecDescr_t vb, vz, vy;
    SP(cusparseCreateDnVec(&vb, n, (void *)b, dt));
    SP(cusparseCreateDnVec(&vz, n, z, dt));
    SP(cusparseCreateDnVec(&vy, n, y, dt));
    cusparseSpSVDescr_t sl, su;
    SP(cusparseSpSV_createDescr(&sl));
    SP(cusparseSpSV_createDescr(&su));
    double ad = 1.0;
    float af = 1.0f;
    const void *alpha = prec ? (const void *)&af : (const void *)&ad;
    size_t bl = 0, bu = 0;
    SP(cusparseSpSV_bufferSize(h, CUSPARSE_OPERATION_NON_TRANSPOSE, alpha, L, vb, vz, dt, CUSPARSE_SPSV_ALG_DEFAULT, sl, &bl));
    SP(cusparseSpSV_bufferSize(h, CUSPARSE_OPERATION_NON_TRANSPOSE, alpha, U, vz, vy, dt, CUSPARSE_SPSV_ALG_DEFAULT, su, &bu));
    void *wl = nullptr, *wu = nullptr;
    CK(cudaMalloc(&wl, bl ? bl : 16));
    CK(cudaMalloc(&wu, bu ? bu : 16));
    Timer tm;
    tm.start();
    SP(cusparseSpSV_analysis(h, CUSPARSE_OPERATION_NON_TRANSPOSE, alpha, L, vb, vz, dt, CUSPARSE_SPSV_ALG_DEFAULT, sl, wl));
    SP(cusparseSpSV_analysis(h, CUSPARSE_OPERATION_NON_TRANSPOSE, alpha, U, vz, vy, dt, CUSPARSE_SPSV_ALG_DEFAULT, su, wu));
    float ta = tm.stop();
    std::vector<float> t(reps > 0 ? reps : 1);
    for (int i = 0; i < (reps > 0 ? reps : 1); ++i) {
        tm.start();
        SP(cusparseSpSV_solve(h, CUSPARSE_OPERATION_NON_TRANSPOSE, alpha, L, vb, vz, dt, CUSPARSE_SPSV_ALG_DEFAULT, sl));
        SP(cusparseSpSV_solve(h, CUSPARSE_OPERATION_NON_TRANSPOSE, alpha, U, vz, vy, dt, CUSPARSE_SPSV_ALG_DEFAULT, su));
        t[i] = tm.stop();
    }
    std::sort(t.begin(), t.end());
    if (ms) { ms[0] = ta; ms[1] = t[t.size() / 2]; }
    cudaFree(wl);
    cudaFree(wu);
    cusparseSpSV_destroyDescr(sl);
    cusparseSpSV_destroyDescr(su);
    cusparseDestroyDnVec(vb);
    cusparseDestroyDnVec(vz);
    cusparseDestroyDnVec(vy);
    cusparseDestroySpMat(L);
    cusparseDestroySpMat(U);
    cusparseDestroy(h);
    return 0;
}
