// api.cu -- the C ABI of libreaspasol_gpu (include/reaspasol_gpu.h): handle, options, entry points,
// host-pointer wrappers and the device-resident matrix object.  Each entry point stands in for
// a cuSPARSE call group of the reference harness (citations in the header).
#include "rsg_internal.cuh"

#include <cstring>
#include <new>

using namespace rsg;

namespace {

template <typename T> struct DevBuf { // RAII for the *_host wrappers
    T *p = nullptr;
    ~DevBuf() { if (p) cudaFree(p); }
    int alloc(size_t n) { return cudaMalloc(&p, sizeof(T) * (n ? n : 1)) == cudaSuccess ? RSG_OK : ((void)cudaGetLastError(), RSG_ALLOC_FAILED); }
    int upload(const T *src, size_t n, cudaStream_t st)
    {
        RSG_TRY(alloc(n));
        if (n) RSG_CUDA_TRY(cudaMemcpyAsync(p, src, sizeof(T) * n, cudaMemcpyHostToDevice, st));
        return RSG_OK;
    }
};

struct DeviceGuard {
    int prev = -1;
    explicit DeviceGuard(int dev) { cudaGetDevice(&prev); if (prev != dev) cudaSetDevice(dev); else prev = -1; }
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

// number of stored entries from HOST row pointers
inline int host_stored(int m, const int *rowptr) { return rowptr[m] - rowptr[0]; }

__global__ void cast_d2f_kernel(size_t n, const double *__restrict__ a, float *__restrict__ b)
{
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        b[i] = (float)a[i];
}

} // namespace

int rsg::check_err_flag(rsg_context *h)
{
    cudaStream_t st = h->stream;
    RSG_CUDA_TRY(cudaMemcpyAsync(h->pinned_i + 63, h->err_flag, sizeof(int), cudaMemcpyDeviceToHost, st));
    RSG_CUDA_TRY(cudaStreamSynchronize(st));
    if (h->pinned_i[63] == 0) return RSG_OK;
    RSG_CUDA_TRY(cudaMemsetAsync(h->err_flag, 0, sizeof(int), st));
    return RSG_DEPENDENCY_TIMEOUT;
}

extern "C" {

int rsg_version(void) { return 100; }

const char *rsg_status_string(int status)
{
    switch (status) {
    case RSG_OK: return "ok";
    case RSG_BAD_ARG: return "bad argument";
    case RSG_CUDA_ERROR: return "CUDA error";
    case RSG_ALLOC_FAILED: return "allocation failed";
    case RSG_NOT_ANALYSED: return "analysis has not been run for this matrix";
    case RSG_ZERO_PIVOT_STRUCTURAL: return "structural zero pivot (missing diagonal entry)";
    case RSG_ZERO_PIVOT_NUMERIC: return "numerical zero pivot";
    case RSG_DEPENDENCY_TIMEOUT: return "dependency wait timed out";
    case RSG_UNSUPPORTED: return "unsupported";
    case RSG_BREAKDOWN: return "Krylov recurrence broke down";
    case RSG_SIZE_LIMIT: return "matrix larger than the fixed scratch of the allocation-free entry point";
    default: return "unknown status";
    }
}

/* debug: the window kernel writes 4 x %globaltimer (ns) per CTA into buf (entry, dependency resolved, data staged, done; bit 63
 * of the last word = the tile reads a peer's entries); buf == NULL switches it off */
int rsg_debug_trace(rsg_handle_t h, unsigned long long *buf, int capacity_ctas)
{
    if (!h || capacity_ctas < 0) return RSG_BAD_ARG;
    h->trace = buf;
    h->trace_ctas = buf ? capacity_ctas : 0;
    return RSG_OK;
}

int rsg_create(rsg_handle_t *handle, int device)
{
    if (!handle) return RSG_BAD_ARG;
    *handle = nullptr;
    int count = 0;
    RSG_CUDA_TRY(cudaGetDeviceCount(&count));
    if (device < 0 || device >= count) return RSG_BAD_ARG;
    DeviceGuard g(device);
    rsg_context *h = new (std::nothrow) rsg_context();
    if (!h) return RSG_ALLOC_FAILED;
    h->device = device;
    cudaDeviceGetAttribute(&h->sm_count, cudaDevAttrMultiProcessorCount, device);
    cudaDeviceGetAttribute(&h->l2_bytes, cudaDevAttrL2CacheSize, device);
    cudaDeviceGetAttribute(&h->max_smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, device);
    h->opts[RSG_OPT_SPMV_CACHE_HINT] = 1;
    if (cudaMalloc(&h->scratch_i, 64 * sizeof(int)) != cudaSuccess || cudaMemset(h->scratch_i, 0, 64 * sizeof(int)) != cudaSuccess ||
        cudaMallocHost(&h->pinned_i, 64 * sizeof(int)) != cudaSuccess) {
        (void)cudaGetLastError();
        if (h->scratch_i) cudaFree(h->scratch_i);
        delete h;
        return RSG_ALLOC_FAILED;
    }
    h->err_flag = h->scratch_i + 60;
    h->spmv_plans.reserve(65);
    h->tri_plans.reserve(9);
    h->transpose_plans.reserve(9);
    *handle = h;
    return RSG_OK;
}

int rsg_destroy(rsg_handle_t h)
{
    if (!h) return RSG_BAD_ARG;
    DeviceGuard g(h->device);
    cudaStreamSynchronize(h->stream);
    for (auto &p : h->spmv_plans) spmv_free_plan(p);
    for (auto &p : h->tri_plans) tri_free_plan(p);
    for (auto &p : h->transpose_plans) transpose_free_plan(p);
    if (h->cub_tmp) cudaFree(h->cub_tmp);
    if (h->krylov_ws) cudaFree(h->krylov_ws);
    if (h->krylov_pin) cudaFreeHost(h->krylov_pin);
    cudaFree(h->scratch_i);
    cudaFreeHost(h->pinned_i);
    delete h;
    return RSG_OK;
}

int rsg_set_stream(rsg_handle_t h, void *cuda_stream)
{
    if (!h) return RSG_BAD_ARG;
    h->stream = static_cast<cudaStream_t>(cuda_stream);
    h->persist_ptr = nullptr;
    return RSG_OK;
}

int rsg_synchronize(rsg_handle_t h)
{
    if (!h) return RSG_BAD_ARG;
    DeviceGuard g(h->device);
    return check_err_flag(h); // synchronises the stream; a device-side wait that gave up surfaces here
}

int rsg_set_option(rsg_handle_t h, int option, int value)
{
    if (!h || option < 0 || option >= RSG_OPT_COUNT_) return RSG_BAD_ARG;
    h->opts[option] = value;
    return RSG_OK;
}

int rsg_get_info(rsg_handle_t h, int info, long long *value)
{
    if (!h || !value) return RSG_BAD_ARG;
    switch (info) {
    case RSG_INFO_KERNEL_LAUNCHES: *value = h->launches; break;
    case RSG_INFO_SPMV_ALGO_USED: *value = h->last_spmv_algo; break;
    case RSG_INFO_SPMV_ROW_BLOCKS: *value = h->last_spmv_blocks; break;
    case RSG_INFO_SPMV_MAX_ROW: *value = h->last_spmv_maxrow; break;
    case RSG_INFO_LEVELS_L: *value = h->last_tri >= 0 ? h->tri_plans[(size_t)h->last_tri].nlevels[0] : 0; break;
    case RSG_INFO_LEVELS_U: *value = h->last_tri >= 0 ? h->tri_plans[(size_t)h->last_tri].nlevels[1] : 0; break;
    case RSG_INFO_SM_COUNT: *value = h->sm_count; break;
    case RSG_INFO_WORKSPACE_BYTES: {
        long long b = (long long)h->cub_tmp_bytes;
        for (auto &p : h->spmv_plans) b += (long long)p.bytes;
        for (auto &p : h->tri_plans) b += (long long)p.bytes;
        *value = b;
        break;
    }
    default: return RSG_BAD_ARG;
    }
    return RSG_OK;
}

int rsg_forget_matrix(rsg_handle_t h, const int *rowptr)
{
    if (!h) return RSG_BAD_ARG;
    DeviceGuard g(h->device);
    cudaStreamSynchronize(h->stream);
    for (size_t i = h->spmv_plans.size(); i-- > 0;)
        if (h->spmv_plans[i].rowptr == rowptr) { spmv_free_plan(h->spmv_plans[i]); h->spmv_plans.erase(h->spmv_plans.begin() + (long)i); }
    for (size_t i = h->tri_plans.size(); i-- > 0;)
        if (h->tri_plans[i].rowptr == rowptr) { tri_free_plan(h->tri_plans[i]); h->tri_plans.erase(h->tri_plans.begin() + (long)i); h->last_tri = -1; }
    for (size_t i = h->transpose_plans.size(); i-- > 0;)
        if (h->transpose_plans[i].rowptr == rowptr) transpose_drop_plan(h, i);
    return RSG_OK;
}

// ---- SpMV ---------------------------------------------------------------------------------
int rsg_spmv_analysis(rsg_handle_t h, int m, int n, int nnz, int base, const int *rowptr, const int *colidx)
{
    if (!h || m < 0 || n < 0 || (base != 0 && base != 1) || !rowptr) return RSG_BAD_ARG;
    DeviceGuard g(h->device);
    SpmvPlan *p = nullptr; // one plan per (stored-value size, vector element size): the tile shapes differ
    RSG_TRY(spmv_get_plan(h, m, n, nnz, base, rowptr, colidx, true, &p, 0x88));
    return spmv_get_plan(h, m, n, nnz, base, rowptr, colidx, true, &p, 0x44); // the mixed plan (0x84) is built on first use
}

int spmv_csr_d(rsg_handle_t h, int m, int n, int nnz, int base, const int *rowptr, const int *colidx,
               const double *values, double alpha, const double *x, double beta, double *y)
{
    if (!h) return RSG_BAD_ARG;
    DeviceGuard g(h->device);
    return spmv_run<double, double>(h, m, n, nnz, base, rowptr, colidx, values, alpha, x, beta, y);
}

int spmv_csr_s(rsg_handle_t h, int m, int n, int nnz, int base, const int *rowptr, const int *colidx,
               const float *values, float alpha, const float *x, float beta, float *y)
{
    if (!h) return RSG_BAD_ARG;
    DeviceGuard g(h->device);
    return spmv_run<float, float>(h, m, n, nnz, base, rowptr, colidx, values, alpha, x, beta, y);
}

int spmv_csr_mixed(rsg_handle_t h, int m, int n, int nnz, int base, const int *rowptr, const int *colidx,
                   const float *values, double alpha, const double *x, double beta, double *y)
{
    if (!h) return RSG_BAD_ARG;
    DeviceGuard g(h->device);
    return spmv_run<float, double>(h, m, n, nnz, base, rowptr, colidx, values, alpha, x, beta, y);
}

// ---- transpose (CSC) path ------------------------------------------------------------------
int rsg_csr2csc_d(rsg_handle_t h, int m, int n, int nnz, int base, const int *rowptr, const int *colidx, const double *values,
                  int *colptr, int *rowidx, double *values_out)
{
    (void)nnz;
    if (!h || m < 0 || n < 0 || (base != 0 && base != 1) || !rowptr || !colptr) return RSG_BAD_ARG;
    DeviceGuard g(h->device);
    return csr2csc_run<double>(h, m, n, base, rowptr, colidx, values, colptr, rowidx, values_out);
}
int rsg_csr2csc_s(rsg_handle_t h, int m, int n, int nnz, int base, const int *rowptr, const int *colidx, const float *values,
                  int *colptr, int *rowidx, float *values_out)
{
    (void)nnz;
    if (!h || m < 0 || n < 0 || (base != 0 && base != 1) || !rowptr || !colptr) return RSG_BAD_ARG;
    DeviceGuard g(h->device);
    return csr2csc_run<float>(h, m, n, base, rowptr, colidx, values, colptr, rowidx, values_out);
}
int spmv_csr_t_d(rsg_handle_t h, int m, int n, int nnz, int base, const int *rowptr, const int *colidx, const double *values,
                 double alpha, const double *x, double beta, double *y)
{
    if (!h || m < 0 || n < 0 || (base != 0 && base != 1) || !rowptr || (n > 0 && !y)) return RSG_BAD_ARG;
    DeviceGuard g(h->device);
    return spmv_transposed_run<double>(h, m, n, nnz, base, rowptr, colidx, values, alpha, x, beta, y);
}
int spmv_csr_t_s(rsg_handle_t h, int m, int n, int nnz, int base, const int *rowptr, const int *colidx, const float *values,
                 float alpha, const float *x, float beta, float *y)
{
    if (!h || m < 0 || n < 0 || (base != 0 && base != 1) || !rowptr || (n > 0 && !y)) return RSG_BAD_ARG;
    DeviceGuard g(h->device);
    return spmv_transposed_run<float>(h, m, n, nnz, base, rowptr, colidx, values, alpha, x, beta, y);
}

// ---- ILU0 / SpTRSV ------------------------------------------------------------------------
int rsg_ilu0_analysis(rsg_handle_t h, int n, int nnz, int base, const int *rowptr, const int *colidx, int *zero_pivot_row)
{
    if (!h || n < 0 || (base != 0 && base != 1) || !rowptr) return RSG_BAD_ARG;
    DeviceGuard g(h->device);
    TriPlan *p = nullptr;
    RSG_TRY(tri_get_plan(h, n, nnz, base, rowptr, colidx, true, &p));
    if (p->unsorted) return RSG_BAD_ARG; // columns must ascend inside every row (see the header)
    if (zero_pivot_row) *zero_pivot_row = p->missing_diag_row;
    if (p->missing_diag_row >= 0) return RSG_ZERO_PIVOT_STRUCTURAL;
    // both patterns in one launch: the L and U dependency chains overlap, and the harness asks for
    // the U levels right afterwards anyway (ilu0.cu:229-252)
    return tri_build_levels(h, p, -1);
}

int rsg_sptrsv_analysis(rsg_handle_t h, int n, int nnz, int base, const int *rowptr, const int *colidx)
{
    if (!h || n < 0 || (base != 0 && base != 1) || !rowptr) return RSG_BAD_ARG;
    DeviceGuard g(h->device);
    TriPlan *p = nullptr;
    RSG_TRY(tri_get_plan(h, n, nnz, base, rowptr, colidx, false, &p));
    if (p->unsorted) return RSG_BAD_ARG;
    if (p->missing_diag_row >= 0) return RSG_ZERO_PIVOT_STRUCTURAL;
    return tri_build_levels(h, p, -1);
}

extern "C++" template <typename T>
int ilu0_factor_impl(rsg_handle_t h, int n, int nnz, int base, const int *rowptr, const int *colidx, T *values, int *zp)
{
    if (!h || n < 0 || (base != 0 && base != 1) || !rowptr || (n > 0 && !values)) return RSG_BAD_ARG;
    DeviceGuard g(h->device);
    TriPlan *p = nullptr;
    RSG_TRY(tri_get_plan(h, n, nnz, base, rowptr, colidx, false, &p));
    if (p->unsorted) return RSG_BAD_ARG;
    return ilu0_run<T>(h, p, values, zp);
}

int ilu0_factor_d(rsg_handle_t h, int n, int nnz, int base, const int *rowptr, const int *colidx, double *values, int *zp)
{
    return ilu0_factor_impl<double>(h, n, nnz, base, rowptr, colidx, values, zp);
}
int ilu0_factor_s(rsg_handle_t h, int n, int nnz, int base, const int *rowptr, const int *colidx, float *values, int *zp)
{
    return ilu0_factor_impl<float>(h, n, nnz, base, rowptr, colidx, values, zp);
}

extern "C++" template <typename T>
int sptrsv_impl(rsg_handle_t h, int n, int nnz, int base, const int *rowptr, const int *colidx, const T *lu, const T *b, T *z, T *y)
{
    if (!h || n < 0 || (base != 0 && base != 1) || !rowptr) return RSG_BAD_ARG;
    DeviceGuard g(h->device);
    TriPlan *p = nullptr;
    RSG_TRY(tri_get_plan(h, n, nnz, base, rowptr, colidx, false, &p));
    if (p->unsorted) return RSG_BAD_ARG;
    return trsv_run<T>(h, p, lu, b, z, y);
}

int sptrsv_lu_d(rsg_handle_t h, int n, int nnz, int base, const int *rowptr, const int *colidx, const double *lu,
                const double *b, double *z, double *y)
{
    return sptrsv_impl<double>(h, n, nnz, base, rowptr, colidx, lu, b, z, y);
}
int sptrsv_lu_s(rsg_handle_t h, int n, int nnz, int base, const int *rowptr, const int *colidx, const float *lu,
                const float *b, float *z, float *y)
{
    return sptrsv_impl<float>(h, n, nnz, base, rowptr, colidx, lu, b, z, y);
}

// ---- introspection ------------------------------------------------------------------------
int rsg_get_levels(rsg_handle_t h, int which, int *nlevels, int *level_ptr, int *level_rows)
{
    if (!h || which < 0 || which > 1 || !nlevels) return RSG_BAD_ARG;
    if (h->last_tri < 0) return RSG_NOT_ANALYSED;
    DeviceGuard g(h->device);
    TriPlan *p = &h->tri_plans[(size_t)h->last_tri];
    if (p->missing_diag_row >= 0 && !p->have[which]) {
        // levels are still well defined without a diagonal: build them on request
    }
    RSG_TRY(tri_build_levels(h, p, which));
    *nlevels = p->nlevels[which];
    if (level_ptr && level_rows) {
        std::memcpy(level_ptr, p->level_ptr_host[which].data(), sizeof(int) * ((size_t)p->nlevels[which] + 1));
        if (p->n > 0)
            RSG_CUDA_TRY(cudaMemcpyAsync(level_rows, p->order[which], sizeof(int) * (size_t)p->n, cudaMemcpyDeviceToHost, h->stream));
        RSG_CUDA_TRY(cudaStreamSynchronize(h->stream));
    }
    return RSG_OK;
}

int rsg_get_row_partition(rsg_handle_t h, int m, int base, const int *rowptr, int ngpu, int balanced, int *row_starts)
{
    if (!h) return RSG_BAD_ARG;
    DeviceGuard g(h->device);
    return row_partition(h, m, base, rowptr, ngpu, balanced, row_starts);
}

// ---- host-pointer wrappers (the harness's cudaMalloc + cudaMemcpy, spmv.cu:91-119,216) -------
extern "C++" template <typename TV, typename TX>
int spmv_host_impl(rsg_handle_t h, int m, int n, int base, const int *rowptr, const int *colidx, const TV *values,
                          TX alpha, const TX *x, TX beta, TX *y)
{
    if (!h || m < 0 || n < 0 || !rowptr) return RSG_BAD_ARG;
    DeviceGuard g(h->device);
    const size_t nnz = (size_t)host_stored(m, rowptr);
    const size_t off = (size_t)(rowptr[0] - base);
    DevBuf<int> d_rp, d_ci;
    DevBuf<TV> d_v;
    DevBuf<TX> d_x, d_y;
    cudaStream_t st = h->stream;
    RSG_TRY(d_rp.upload(rowptr, (size_t)m + 1, st));
    // entries are uploaded from position 0 so that row pointers index them directly
    RSG_TRY(d_ci.upload(colidx, off + nnz, st));
    RSG_TRY(d_v.upload(values, off + nnz, st));
    RSG_TRY(d_x.upload(x, (size_t)n, st));
    RSG_TRY(d_y.alloc((size_t)m));
    if (beta != (TX)0 && m) RSG_CUDA_TRY(cudaMemcpyAsync(d_y.p, y, sizeof(TX) * (size_t)m, cudaMemcpyHostToDevice, st));
    int rc = spmv_run<TV, TX>(h, m, n, (int)nnz, base, d_rp.p, d_ci.p, d_v.p, alpha, d_x.p, beta, d_y.p);
    if (rc == RSG_OK && m) {
        if (cudaMemcpyAsync(y, d_y.p, sizeof(TX) * (size_t)m, cudaMemcpyDeviceToHost, st) != cudaSuccess) rc = RSG_CUDA_ERROR;
    }
    if (cudaStreamSynchronize(st) != cudaSuccess && rc == RSG_OK) rc = RSG_CUDA_ERROR;
    // the plan was keyed on temporary device arrays: drop it
    for (size_t i = 0; i < h->spmv_plans.size(); ++i)
        if (h->spmv_plans[i].rowptr == d_rp.p) { spmv_free_plan(h->spmv_plans[i]); h->spmv_plans.erase(h->spmv_plans.begin() + (long)i); break; }
    h->persist_ptr = nullptr;
    return rc;
}

int spmv_csr_d_host(rsg_handle_t h, int m, int n, int nnz, int base, const int *rowptr, const int *colidx,
                    const double *values, double alpha, const double *x, double beta, double *y)
{
    (void)nnz;
    return spmv_host_impl<double, double>(h, m, n, base, rowptr, colidx, values, alpha, x, beta, y);
}
int spmv_csr_s_host(rsg_handle_t h, int m, int n, int nnz, int base, const int *rowptr, const int *colidx,
                    const float *values, float alpha, const float *x, float beta, float *y)
{
    (void)nnz;
    return spmv_host_impl<float, float>(h, m, n, base, rowptr, colidx, values, alpha, x, beta, y);
}
int spmv_csr_mixed_host(rsg_handle_t h, int m, int n, int nnz, int base, const int *rowptr, const int *colidx,
                        const float *values, double alpha, const double *x, double beta, double *y)
{
    (void)nnz;
    return spmv_host_impl<float, double>(h, m, n, base, rowptr, colidx, values, alpha, x, beta, y);
}

static void drop_tri_plan(rsg_handle_t h, const int *rowptr_dev)
{
    for (size_t i = 0; i < h->tri_plans.size(); ++i)
        if (h->tri_plans[i].rowptr == rowptr_dev) {
            tri_free_plan(h->tri_plans[i]);
            h->tri_plans.erase(h->tri_plans.begin() + (long)i);
            h->last_tri = -1;
            break;
        }
}

extern "C++" template <typename T>
int ilu0_host_impl(rsg_handle_t h, int n, int base, const int *rowptr, const int *colidx, T *values, int *zp)
{
    if (!h || n < 0 || !rowptr) return RSG_BAD_ARG;
    DeviceGuard g(h->device);
    const size_t nnz = (size_t)host_stored(n, rowptr) + (size_t)(rowptr[0] - base);
    DevBuf<int> d_rp, d_ci;
    DevBuf<T> d_v;
    cudaStream_t st = h->stream;
    RSG_TRY(d_rp.upload(rowptr, (size_t)n + 1, st));
    RSG_TRY(d_ci.upload(colidx, nnz, st));
    RSG_TRY(d_v.upload(values, nnz, st));
    int dummy = -1;
    int rc = ilu0_factor_impl<T>(h, n, (int)nnz, base, d_rp.p, d_ci.p, d_v.p, zp ? zp : &dummy);
    if ((rc == RSG_OK || rc == RSG_ZERO_PIVOT_NUMERIC) && nnz) {
        if (cudaMemcpyAsync(values, d_v.p, sizeof(T) * nnz, cudaMemcpyDeviceToHost, st) != cudaSuccess) rc = RSG_CUDA_ERROR;
    }
    if (cudaStreamSynchronize(st) != cudaSuccess && rc == RSG_OK) rc = RSG_CUDA_ERROR;
    drop_tri_plan(h, d_rp.p);
    return rc;
}

int ilu0_factor_d_host(rsg_handle_t h, int n, int nnz, int base, const int *rowptr, const int *colidx, double *values, int *zp)
{
    (void)nnz;
    return ilu0_host_impl<double>(h, n, base, rowptr, colidx, values, zp);
}
int ilu0_factor_s_host(rsg_handle_t h, int n, int nnz, int base, const int *rowptr, const int *colidx, float *values, int *zp)
{
    (void)nnz;
    return ilu0_host_impl<float>(h, n, base, rowptr, colidx, values, zp);
}

extern "C++" template <typename T>
int sptrsv_host_impl(rsg_handle_t h, int n, int base, const int *rowptr, const int *colidx, const T *lu, const T *b, T *y)
{
    if (!h || n < 0 || !rowptr) return RSG_BAD_ARG;
    DeviceGuard g(h->device);
    const size_t nnz = (size_t)host_stored(n, rowptr) + (size_t)(rowptr[0] - base);
    DevBuf<int> d_rp, d_ci;
    DevBuf<T> d_v, d_b, d_z, d_y;
    cudaStream_t st = h->stream;
    RSG_TRY(d_rp.upload(rowptr, (size_t)n + 1, st));
    RSG_TRY(d_ci.upload(colidx, nnz, st));
    RSG_TRY(d_v.upload(lu, nnz, st));
    RSG_TRY(d_b.upload(b, (size_t)n, st));
    RSG_TRY(d_z.alloc((size_t)n));
    RSG_TRY(d_y.alloc((size_t)n));
    int rc = sptrsv_impl<T>(h, n, (int)nnz, base, d_rp.p, d_ci.p, d_v.p, d_b.p, d_z.p, d_y.p);
    if (rc == RSG_OK && n) {
        if (cudaMemcpyAsync(y, d_y.p, sizeof(T) * (size_t)n, cudaMemcpyDeviceToHost, st) != cudaSuccess) rc = RSG_CUDA_ERROR;
    }
    if (cudaStreamSynchronize(st) != cudaSuccess && rc == RSG_OK) rc = RSG_CUDA_ERROR;
    drop_tri_plan(h, d_rp.p);
    return rc;
}

int sptrsv_lu_d_host(rsg_handle_t h, int n, int nnz, int base, const int *rowptr, const int *colidx, const double *lu,
                     const double *b, double *y)
{
    (void)nnz;
    return sptrsv_host_impl<double>(h, n, base, rowptr, colidx, lu, b, y);
}
int sptrsv_lu_s_host(rsg_handle_t h, int n, int nnz, int base, const int *rowptr, const int *colidx, const float *lu,
                     const float *b, float *y)
{
    (void)nnz;
    return sptrsv_host_impl<float>(h, n, base, rowptr, colidx, lu, b, y);
}

// ---- device-resident matrix ----------------------------------------------------------------
int rsg_matrix_upload(rsg_handle_t h, char precision, int m, int n, int nnz, int base, const int *rowptr_host,
                      const int *colidx_host, const double *values_host, rsg_matrix_t *matrix)
{
    (void)nnz;
    if (!h || !matrix || m < 0 || n < 0 || !rowptr_host || (precision != 'd' && precision != 's' && precision != 'm'))
        return RSG_BAD_ARG;
    *matrix = nullptr;
    DeviceGuard g(h->device);
    const size_t cnt = (size_t)host_stored(m, rowptr_host) + (size_t)(rowptr_host[0] - base);
    rsg_matrix *a = new (std::nothrow) rsg_matrix();
    if (!a) return RSG_ALLOC_FAILED;
    std::memset(a, 0, sizeof(*a));
    a->precision = precision; a->m = m; a->n = n; a->nnz = (int)cnt; a->base = base;
    const size_t vs = precision == 'd' ? 8 : 4, xs = precision == 's' ? 4 : 8;
    cudaStream_t st = h->stream;
    bool ok = cudaMalloc(&a->rowptr, sizeof(int) * ((size_t)m + 1)) == cudaSuccess &&
              cudaMalloc(&a->colidx, sizeof(int) * (cnt ? cnt : 1)) == cudaSuccess &&
              cudaMalloc(&a->values, vs * (cnt ? cnt : 1)) == cudaSuccess &&
              cudaMalloc(&a->x, xs * (size_t)(n ? n : 1)) == cudaSuccess &&
              cudaMalloc(&a->y, xs * (size_t)(m ? m : 1)) == cudaSuccess &&
              cudaMallocHost(&a->pin_x, xs * (size_t)(n ? n : 1)) == cudaSuccess &&
              cudaMallocHost(&a->pin_y, xs * (size_t)(m ? m : 1)) == cudaSuccess;
    if (!ok) { (void)cudaGetLastError(); rsg_matrix_free(h, a); return RSG_ALLOC_FAILED; }
    int rc = RSG_OK;
    if (cudaMemcpyAsync(a->rowptr, rowptr_host, sizeof(int) * ((size_t)m + 1), cudaMemcpyHostToDevice, st) != cudaSuccess) rc = RSG_CUDA_ERROR;
    if (cnt && cudaMemcpyAsync(a->colidx, colidx_host, sizeof(int) * cnt, cudaMemcpyHostToDevice, st) != cudaSuccess) rc = RSG_CUDA_ERROR;
    if (cnt && rc == RSG_OK) {
        if (precision == 'd') {
            if (cudaMemcpyAsync(a->values, values_host, 8 * cnt, cudaMemcpyHostToDevice, st) != cudaSuccess) rc = RSG_CUDA_ERROR;
        } else { // the harness casts on the host (spmv.cu:63-67); here the cast is fused into the upload
            double *tmp = nullptr;
            if (cudaMalloc(&tmp, 8 * cnt) != cudaSuccess) rc = RSG_ALLOC_FAILED;
            else {
                if (cudaMemcpyAsync(tmp, values_host, 8 * cnt, cudaMemcpyHostToDevice, st) != cudaSuccess) rc = RSG_CUDA_ERROR;
                cast_d2f_kernel<<<h->sm_count * 4, 256, 0, st>>>(cnt, tmp, static_cast<float *>(a->values));
                h->launches++;
                cudaStreamSynchronize(st);
                cudaFree(tmp);
            }
        }
    }
    if (cudaStreamSynchronize(st) != cudaSuccess && rc == RSG_OK) rc = RSG_CUDA_ERROR;
    if (rc != RSG_OK) { (void)cudaGetLastError(); rsg_matrix_free(h, a); return rc; }
    // Row chunks of the pipelined host-vector product: entry-balanced cuts; chunk c needs x[0 .. hi_c) where
    // hi_c = 1 + the largest zero-based column any row of chunks 0..c references (one pass over colidx).
    int want = h->opts[RSG_OPT_HOST_CHUNKS] > 0 ? h->opts[RSG_OPT_HOST_CHUNKS] : 4; // measured best on config 2 (tools/e2e_breakdown.py)
    want = std::min(want, (int)rsg_matrix::kMaxChunks);
    if ((size_t)m * xs < ((size_t)1 << 20) || cnt == 0) want = 1; // small vectors: one piece, latency wins
    a->nchunk = 1;
    a->chunk_row[0] = 0; a->chunk_row[1] = m;
    a->chunk_x[0] = 0; a->chunk_x[1] = n;
    if (want > 1) {
        const long long first = rowptr_host[0], total = (long long)rowptr_host[m] - first;
        int nc = 0;
        for (int c = 1; c <= want; ++c) {
            int r = m;
            if (c < want) // cut on a multiple of 4 rows: every chunk's y slice stays 16/32-byte aligned
                r = (int)(std::lower_bound(rowptr_host, rowptr_host + m + 1, (int)(first + total * c / want)) - rowptr_host) & ~3;
            r = std::min(std::max(r, a->chunk_row[nc]), m);
            if (r > a->chunk_row[nc]) a->chunk_row[++nc] = r;
        }
        a->nchunk = nc;
        int hi = 0;
        for (int c = 0; c < nc; ++c) {
            const int *ci = colidx_host + (rowptr_host[a->chunk_row[c]] - base), *ce = colidx_host + (rowptr_host[a->chunk_row[c + 1]] - base);
            int mx = -1;
            for (; ci < ce; ++ci) mx = std::max(mx, *ci);
            hi = std::max(hi, std::min(mx - base + 1, n));
            a->chunk_x[c + 1] = c + 1 == nc ? n : hi;
        }
    }
    bool ev_ok = cudaStreamCreateWithFlags(&a->s_in, cudaStreamNonBlocking) == cudaSuccess &&
                 cudaStreamCreateWithFlags(&a->s_out, cudaStreamNonBlocking) == cudaSuccess &&
                 cudaStreamCreateWithFlags(&a->s_main, cudaStreamNonBlocking) == cudaSuccess &&
                 cudaEventCreateWithFlags(&a->ev_start, cudaEventDisableTiming) == cudaSuccess &&
                 cudaEventCreateWithFlags(&a->ev_pre, cudaEventDisableTiming) == cudaSuccess &&
                 cudaEventCreateWithFlags(&a->ev_join_in, cudaEventDisableTiming) == cudaSuccess &&
                 cudaEventCreateWithFlags(&a->ev_join_out, cudaEventDisableTiming) == cudaSuccess;
    for (int c = 0; ev_ok && c < a->nchunk; ++c)
        ev_ok = cudaEventCreateWithFlags(&a->ev_in[c], cudaEventDisableTiming) == cudaSuccess &&
                cudaEventCreateWithFlags(&a->ev_mul[c], cudaEventDisableTiming) == cudaSuccess;
    if (!ev_ok) { (void)cudaGetLastError(); rsg_matrix_free(h, a); return RSG_CUDA_ERROR; }
    // plans of the whole matrix and of every chunk view now, so that no product pays for an analysis
    SpmvPlan *pl = nullptr;
    const int pkey = (int)vs | ((int)xs << 4);
    rc = spmv_get_plan(h, m, n, a->nnz, base, a->rowptr, a->colidx, false, &pl, pkey);
    for (int c = 0; rc == RSG_OK && a->nchunk > 1 && c < a->nchunk; ++c)
        rc = spmv_get_plan(h, a->chunk_row[c + 1] - a->chunk_row[c], n, a->nnz, base, a->rowptr + a->chunk_row[c], a->colidx,
                           false, &pl, pkey);
    if (rc != RSG_OK) { rsg_matrix_free(h, a); return rc; }
    *matrix = a;
    return RSG_OK;
}

extern "C++" template <typename TV, typename TX>
int matrix_rows_spmv(rsg_handle_t h, rsg_matrix_t a, int r0, int r1, double alpha, double beta, void *y)
{
    return spmv_run<TV, TX>(h, r1 - r0, a->n, a->nnz, a->base, a->rowptr + r0, a->colidx, static_cast<const TV *>(a->values),
                            (TX)alpha, static_cast<const TX *>(a->x), (TX)beta, static_cast<TX *>(y) + r0);
}
// y: where row 0 of the product goes (device memory, or page-locked host memory the device can address)
static int matrix_rows_spmv_any(rsg_handle_t h, rsg_matrix_t a, int r0, int r1, double alpha, double beta, void *y)
{
    if (a->precision == 'd') return matrix_rows_spmv<double, double>(h, a, r0, r1, alpha, beta, y);
    if (a->precision == 's') return matrix_rows_spmv<float, float>(h, a, r0, r1, alpha, beta, y);
    return matrix_rows_spmv<float, double>(h, a, r0, r1, alpha, beta, y);
}

int rsg_matrix_spmv_host(rsg_handle_t h, rsg_matrix_t a, double alpha, const void *x_host, double beta, void *y_host)
{
    if (!h || !a || !y_host || (a->n > 0 && !x_host)) return RSG_BAD_ARG;
    DeviceGuard g(h->device);
    cudaStream_t st = h->stream;
    const size_t xs = a->precision == 's' ? 4 : 8;
    // pageable -> pinned staging -> device, like the harness's cudaMemcpy but asynchronous; a caller
    // buffer that is already page-locked (cudaMallocHost / cudaHostRegister) is copied from directly
    auto pinned = [](const void *p) {
        cudaPointerAttributes at;
        if (cudaPointerGetAttributes(&at, p) != cudaSuccess) { (void)cudaGetLastError(); return false; }
        return at.type == cudaMemoryTypeHost;
    };
    const bool x_pinned = a->n > 0 && pinned(x_host), y_pinned = pinned(y_host);
    if (!x_pinned) std::memcpy(a->pin_x, x_host, xs * (size_t)a->n);
    const char *xsrc = static_cast<const char *>(x_pinned ? x_host : a->pin_x);
    char *ydst = static_cast<char *>(y_pinned ? y_host : a->pin_y);
    if (beta != 0.0 || a->nchunk <= 1) { // one piece: x in, product, y out on the handle's stream
        RSG_CUDA_TRY(cudaMemcpyAsync(a->x, xsrc, xs * (size_t)a->n, cudaMemcpyHostToDevice, st));
        if (beta != 0.0) {
            if (!y_pinned) std::memcpy(a->pin_y, y_host, xs * (size_t)a->m);
            RSG_CUDA_TRY(cudaMemcpyAsync(a->y, ydst, xs * (size_t)a->m, cudaMemcpyHostToDevice, st));
        }
        RSG_TRY(matrix_rows_spmv_any(h, a, 0, a->m, alpha, beta, a->y));
        RSG_CUDA_TRY(cudaMemcpyAsync(ydst, a->y, xs * (size_t)a->m, cudaMemcpyDeviceToHost, st));
        RSG_CUDA_TRY(cudaStreamSynchronize(st));
    } else {
        // s_in carries x upward piece by piece; s_main multiplies chunk c once its prefix of x has landed; s_out copies
        // chunk c's slice of y back while chunk c+1 is multiplied (HOST_Y_DIRECT = 1: the kernels store straight into the
        // page-locked host buffer instead -- measured slower, tools/e2e_breakdown.py).  The ~30 stream calls of one step
        // are captured into a CUDA graph the second time the same buffers and alpha come by; later steps are one launch.
        const bool direct = h->opts[RSG_OPT_HOST_Y_DIRECT] == 1;
        void *ydev = a->y;
        if (direct && cudaHostGetDevicePointer(&ydev, ydst, 0) != cudaSuccess) { (void)cudaGetLastError(); return RSG_CUDA_ERROR; }
        auto issue = [&]() -> int { // fork from s_main, join back into it
            RSG_CUDA_TRY(cudaEventRecord(a->ev_start, a->s_main));
            RSG_CUDA_TRY(cudaStreamWaitEvent(a->s_in, a->ev_start, 0));
            RSG_CUDA_TRY(cudaStreamWaitEvent(a->s_out, a->ev_start, 0));
            for (int c = 0; c < a->nchunk; ++c) {
                const size_t x0 = (size_t)a->chunk_x[c], x1 = (size_t)a->chunk_x[c + 1];
                if (x1 > x0) RSG_CUDA_TRY(cudaMemcpyAsync(static_cast<char *>(a->x) + xs * x0, xsrc + xs * x0, xs * (x1 - x0), cudaMemcpyHostToDevice, a->s_in));
                RSG_CUDA_TRY(cudaEventRecord(a->ev_in[c], a->s_in));
            }
            cudaStream_t keep = h->stream;
            h->stream = a->s_main; // the SpMV launches go where the pipeline lives
            int rc = RSG_OK;
            for (int c = 0; c < a->nchunk && rc == RSG_OK; ++c) {
                const int r0 = a->chunk_row[c], r1 = a->chunk_row[c + 1];
                if (cudaStreamWaitEvent(a->s_main, a->ev_in[c], 0) != cudaSuccess) { rc = RSG_CUDA_ERROR; break; }
                if (r1 > r0) rc = matrix_rows_spmv_any(h, a, r0, r1, alpha, 0.0, ydev);
                if (direct || rc != RSG_OK) continue;
                if (cudaEventRecord(a->ev_mul[c], a->s_main) != cudaSuccess || cudaStreamWaitEvent(a->s_out, a->ev_mul[c], 0) != cudaSuccess ||
                    (r1 > r0 && cudaMemcpyAsync(ydst + xs * (size_t)r0, static_cast<char *>(a->y) + xs * (size_t)r0, xs * (size_t)(r1 - r0),
                                                cudaMemcpyDeviceToHost, a->s_out) != cudaSuccess))
                    rc = RSG_CUDA_ERROR;
            }
            h->stream = keep;
            RSG_TRY(rc);
            RSG_CUDA_TRY(cudaEventRecord(a->ev_join_in, a->s_in));
            RSG_CUDA_TRY(cudaStreamWaitEvent(a->s_main, a->ev_join_in, 0));
            RSG_CUDA_TRY(cudaEventRecord(a->ev_join_out, a->s_out));
            RSG_CUDA_TRY(cudaStreamWaitEvent(a->s_main, a->ev_join_out, 0));
            return RSG_OK;
        };
        // order the step after whatever the handle's stream still has in flight
        RSG_CUDA_TRY(cudaEventRecord(a->ev_pre, st));
        RSG_CUDA_TRY(cudaStreamWaitEvent(a->s_main, a->ev_pre, 0));
        int sig = (direct ? 1 : 0) + a->nchunk * 2;
        for (int o = 0; o < RSG_OPT_COUNT_; ++o) sig = sig * 31 + h->opts[o];
        sig = sig * 31 + (int)spmv_plans_freed(); // a freed plan may be one the captured kernels point into
        const bool want_graph = h->opts[RSG_OPT_HOST_GRAPH] != 2;
        const bool same_as_graph = a->graph && a->g_x == xsrc && a->g_y == ydst && a->g_alpha == alpha && a->g_sig == sig;
        if (want_graph && !same_as_graph && a->seen_x == xsrc && a->seen_y == ydst && a->seen_alpha == alpha && a->seen_sig == sig) {
            // second visit of these buffers: capture (every kernel involved has been launched once, plans exist)
            if (a->graph) { cudaGraphExecDestroy(a->graph); a->graph = nullptr; }
            cudaGraph_t graph = nullptr;
            const long long before = h->launches;
            if (cudaStreamBeginCapture(a->s_main, cudaStreamCaptureModeRelaxed) == cudaSuccess) {
                const int rc = issue();
                const cudaError_t e = cudaStreamEndCapture(a->s_main, &graph);
                if (rc == RSG_OK && e == cudaSuccess && graph && cudaGraphInstantiate(&a->graph, graph, 0) == cudaSuccess) {
                    a->g_x = xsrc; a->g_y = ydst; a->g_alpha = alpha; a->g_sig = sig;
                    a->g_kernels = (int)(h->launches - before);
                } else {
                    (void)cudaGetLastError();
                    a->graph = nullptr;
                }
                if (graph) cudaGraphDestroy(graph);
            } else {
                (void)cudaGetLastError();
            }
            h->launches = before; // captured, not run
        }
        a->seen_x = xsrc; a->seen_y = ydst; a->seen_alpha = alpha; a->seen_sig = sig;
        if (want_graph && a->graph && a->g_x == xsrc && a->g_y == ydst && a->g_alpha == alpha && a->g_sig == sig) {
            RSG_CUDA_TRY(cudaGraphLaunch(a->graph, a->s_main));
            h->launches += a->g_kernels;
        } else {
            RSG_TRY(issue());
        }
        if (cudaStreamSynchronize(a->s_main) != cudaSuccess) { (void)cudaGetLastError(); return RSG_CUDA_ERROR; }
    }
    if (!y_pinned) std::memcpy(y_host, a->pin_y, xs * (size_t)a->m);
    return RSG_OK;
}

int rsg_matrix_free(rsg_handle_t h, rsg_matrix_t a)
{
    if (!h || !a) return RSG_BAD_ARG;
    DeviceGuard g(h->device);
    for (size_t i = h->spmv_plans.size(); i-- > 0;) // the matrix's plan and those of its chunk views
        if (a->rowptr && h->spmv_plans[i].rowptr >= a->rowptr && h->spmv_plans[i].rowptr <= a->rowptr + a->m) {
            spmv_free_plan(h->spmv_plans[i]);
            h->spmv_plans.erase(h->spmv_plans.begin() + (long)i);
        }
    cudaFree(a->rowptr); cudaFree(a->colidx); cudaFree(a->values); cudaFree(a->x); cudaFree(a->y);
    if (a->pin_x) cudaFreeHost(a->pin_x);
    if (a->pin_y) cudaFreeHost(a->pin_y);
    if (a->s_in) cudaStreamDestroy(a->s_in);
    if (a->s_out) cudaStreamDestroy(a->s_out);
    if (a->s_main) cudaStreamDestroy(a->s_main);
    if (a->graph) cudaGraphExecDestroy(a->graph);
    if (a->ev_start) cudaEventDestroy(a->ev_start);
    if (a->ev_pre) cudaEventDestroy(a->ev_pre);
    if (a->ev_join_in) cudaEventDestroy(a->ev_join_in);
    if (a->ev_join_out) cudaEventDestroy(a->ev_join_out);
    for (int c = 0; c < rsg_matrix::kMaxChunks; ++c) {
        if (a->ev_in[c]) cudaEventDestroy(a->ev_in[c]);
        if (a->ev_mul[c]) cudaEventDestroy(a->ev_mul[c]);
    }
    delete a;
    return RSG_OK;
}

} // extern "C"
