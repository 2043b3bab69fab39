// transpose.cu -- CSR -> CSC on the device and y = alpha*A^T*x + beta*y.
// The reference gets a CSC by asking its loader for the transposed matrix (transpose = 1,
// /root/reference/ReadMatrixMarket/loadMatrixMarket.cpp:79-81,124; used by test_superLU_MT.c:85): the same
// arrays are produced here from device-resident CSR without going back to the file -- a STABLE radix sort of the
// column indices (CUB) carrying the entry positions, so inside a column the entries keep their row order, which is
// what the loader's per-row sort of the transposed matrix yields (bit-identical for matrices without duplicate
// coordinates; tests/test_gpu_transpose.py).  The transposed product then runs the ordinary SpMV kernels on the
// transposed pattern (deterministic, no atomics); the pattern and the permutation are cached per handle.  The values
// may have changed between calls (e.g. after a factorization), so by default they are read through the permutation
// inside the product (spmv_t_fused_kernel) or, for patterns with long columns, re-gathered into a transposed copy on
// every call -- unless the caller sets RSG_OPT_TRANSPOSE_STATIC = 1 and so promises they do not change: then the copy
// is made once and the product is the tile kernel's.
#include <cub/device/device_radix_sort.cuh>

#include "rsg_internal.cuh"

namespace rsg {

__global__ void tr_iota_cols_kernel(int nnz, int first, int base, const int *__restrict__ colidx, int *__restrict__ keys, int *__restrict__ pos)
{
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= nnz) return;
    keys[k] = colidx[first + k] - base;
    pos[k] = k;
}

// colptr[c] = base + number of entries with column < c  (binary search in the sorted keys)
__global__ void tr_colptr_kernel(int n, int nnz, int base, const int *__restrict__ sorted_cols, int *__restrict__ colptr)
{
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c > n) return;
    int lo = 0, hi = nnz;
    while (lo < hi) {
        const int mid = lo + ((hi - lo) >> 1);
        if (sorted_cols[mid] < c) lo = mid + 1; else hi = mid;
    }
    colptr[c] = lo + base;
}

// rowidx[k] = base + row of the entry that landed at k  (binary search in rowptr)
__global__ void tr_rowidx_kernel(int m, int nnz, int base, const int *__restrict__ rowptr, const int *__restrict__ perm,
                                 int *__restrict__ rowidx)
{
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= nnz) return;
    const int e = perm[k] + rowptr[0]; // position in the caller's numbering (rowptr values carry the base)
    int lo = 0, hi = m; // largest r with rowptr[r] <= e
    while (lo < hi) {
        const int mid = lo + ((hi - lo + 1) >> 1);
        if (rowptr[mid] <= e) lo = mid; else hi = mid - 1;
    }
    rowidx[k] = lo + base;
}

template <typename T>
__global__ void tr_gather_kernel(int nnz, const int *__restrict__ perm, const T *__restrict__ values, T *__restrict__ out)
{
    for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < nnz; k += gridDim.x * blockDim.x) out[k] = values[perm[k]];
}

// y = alpha*A^T*x + beta*y with the values fetched THROUGH the permutation, so no transposed copy of them is written
// and read back: per stored entry 4 B (row index) + 4 B (perm) + 8/4 B (value, each sector shared by the neighbouring
// columns) instead of the 20 B gather pass + 12 B product.  G lanes per column of A, straight from global memory;
// fixed summation order (lane-strided partial sums, xor tree), so the result is deterministic.  Only for patterns
// whose longest column is short (a lane walks its share serially): the caller falls back to gather + tile kernel.
template <typename T, int G, int B>
__global__ void __launch_bounds__(256)
spmv_t_fused_kernel(int ncols, int base, const int *__restrict__ colptr, const int *__restrict__ rowidx, const int *__restrict__ perm,
                    const T *__restrict__ values, const T *__restrict__ x, T *__restrict__ y, T alpha, T beta)
{
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    const int c = t / G, lane = t % G;
    T s = 0;
    if (c < ncols) {
        const int e = colptr[c + 1] - base;
        int k = colptr[c] - base + lane;
        for (; k + (B - 1) * G < e; k += B * G) { // B independent entry chains in flight per lane
            int pp[B], rr[B];
            T aa[B], xx[B];
#pragma unroll
            for (int j = 0; j < B; ++j) { pp[j] = perm[k + j * G]; rr[j] = rowidx[k + j * G] - base; }
#pragma unroll
            for (int j = 0; j < B; ++j) { aa[j] = __ldg(values + pp[j]); xx[j] = __ldg(x + rr[j]); }
#pragma unroll
            for (int j = 0; j < B; ++j) s += aa[j] * xx[j];
        }
        if (k < e) { // the tail as one predicated batch, so its loads overlap as well
            int pp[B], rr[B];
            T aa[B], xx[B];
#pragma unroll
            for (int j = 0; j < B; ++j) {
                const bool ok = k + j * G < e;
                pp[j] = ok ? perm[k + j * G] : 0;
                rr[j] = ok ? rowidx[k + j * G] - base : 0;
            }
#pragma unroll
            for (int j = 0; j < B; ++j) {
                const bool ok = k + j * G < e;
                aa[j] = ok ? __ldg(values + pp[j]) : (T)0;
                xx[j] = ok ? __ldg(x + rr[j]) : (T)0;
            }
#pragma unroll
            for (int j = 0; j < B; ++j)
                if (k + j * G < e) s += aa[j] * xx[j];
        }
    }
#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (c < ncols && lane == 0) y[c] = (beta == (T)0) ? alpha * s : alpha * s + beta * y[c];
}

void transpose_free_plan(TransposePlan &p)
{
    cudaFree(p.colptr); cudaFree(p.rowidx); cudaFree(p.perm); cudaFree(p.values_t);
    p = TransposePlan();
}

// drops plan i together with the SpMV plan that was keyed on its (about to be recycled) device arrays
void transpose_drop_plan(rsg_context *h, size_t i)
{
    const int *key = h->transpose_plans[i].colptr;
    for (size_t j = h->spmv_plans.size(); j-- > 0;)
        if (h->spmv_plans[j].rowptr == key) { spmv_free_plan(h->spmv_plans[j]); h->spmv_plans.erase(h->spmv_plans.begin() + (long)j); }
    transpose_free_plan(h->transpose_plans[i]);
    h->transpose_plans.erase(h->transpose_plans.begin() + (long)i);
}

// colptr (n+1), rowidx (nnz), perm (nnz): all device, caller-provided.  perm[k] = zero-based position (relative to
// the first stored entry) of the entry that lands at k.
static int build_transpose(rsg_context *h, int m, int n, int nnz, int first, int base, const int *rowptr, const int *colidx,
                           int *colptr, int *rowidx, int *perm)
{
    cudaStream_t st = h->stream;
    if (nnz == 0) {
        tr_colptr_kernel<<<(n + 1 + 255) / 256, 256, 0, st>>>(n, 0, base, nullptr, colptr);
        h->launches++;
        RSG_CUDA_TRY(cudaGetLastError());
        return RSG_OK;
    }
    int *keys = nullptr, *pos = nullptr, *skeys = nullptr;
    RSG_CUDA_TRY(cudaMalloc(&keys, sizeof(int) * (size_t)nnz * 3));
    pos = keys + nnz;
    skeys = pos + nnz;
    tr_iota_cols_kernel<<<(nnz + 255) / 256, 256, 0, st>>>(nnz, first, base, colidx, keys, pos);
    int end_bit = 1;
    while (end_bit < 31 && (1LL << end_bit) < (long long)n) ++end_bit;
    size_t tmp = 0;
    int rc = RSG_OK;
    if (cub::DeviceRadixSort::SortPairs(nullptr, tmp, keys, skeys, pos, perm, nnz, 0, end_bit, st) != cudaSuccess) rc = RSG_CUDA_ERROR;
    if (rc == RSG_OK) rc = ensure_cub_tmp(h, tmp);
    if (rc == RSG_OK && cub::DeviceRadixSort::SortPairs(h->cub_tmp, tmp, keys, skeys, pos, perm, nnz, 0, end_bit, st) != cudaSuccess)
        rc = RSG_CUDA_ERROR;
    if (rc == RSG_OK) {
        tr_colptr_kernel<<<(n + 1 + 255) / 256, 256, 0, st>>>(n, nnz, base, skeys, colptr);
        tr_rowidx_kernel<<<(nnz + 255) / 256, 256, 0, st>>>(m, nnz, base, rowptr, perm, rowidx);
        h->launches += 6;
        if (cudaGetLastError() != cudaSuccess) rc = RSG_CUDA_ERROR;
    }
    if (cudaStreamSynchronize(st) != cudaSuccess && rc == RSG_OK) rc = RSG_CUDA_ERROR;
    cudaFree(keys);
    if (rc != RSG_OK) (void)cudaGetLastError();
    return rc;
}

// first stored entry (zero-based) and stored count of a (possibly row-slab) CSR, read from the device
static int stored_range(rsg_context *h, int m, int base, const int *rowptr, int *first, int *nnz)
{
    RSG_CUDA_TRY(cudaMemcpyAsync(h->pinned_i + 16, rowptr, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    RSG_CUDA_TRY(cudaMemcpyAsync(h->pinned_i + 17, rowptr + m, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    RSG_CUDA_TRY(cudaStreamSynchronize(h->stream));
    *first = h->pinned_i[16] - base;
    *nnz = h->pinned_i[17] - h->pinned_i[16];
    return (*first < 0 || *nnz < 0) ? RSG_BAD_ARG : RSG_OK;
}

template <typename T>
int csr2csc_run(rsg_context *h, int m, int n, int base, const int *rowptr, const int *colidx, const T *values, int *colptr,
                int *rowidx, T *values_out)
{
    int first = 0, nnz = 0;
    RSG_TRY(stored_range(h, m, base, rowptr, &first, &nnz));
    int *perm = nullptr;
    RSG_CUDA_TRY(cudaMalloc(&perm, sizeof(int) * (size_t)std::max(nnz, 1)));
    int rc = build_transpose(h, m, n, nnz, first, base, rowptr, colidx, colptr, rowidx, perm);
    if (rc == RSG_OK && nnz > 0 && values && values_out) {
        tr_gather_kernel<T><<<std::min((nnz + 255) / 256, h->sm_count * 16), 256, 0, h->stream>>>(nnz, perm, values + first, values_out);
        h->launches++;
        if (cudaStreamSynchronize(h->stream) != cudaSuccess) rc = RSG_CUDA_ERROR;
    }
    cudaFree(perm);
    return rc;
}

constexpr int kFusedMaxColumn = 2048; // longest column the fused kernel takes (a lane walks column/G entries serially)

template <typename T>
int spmv_transposed_run(rsg_context *h, int m, int n, int nnz_hint, int base, const int *rowptr, const int *colidx, const T *values,
                        T alpha, const T *x, T beta, T *y)
{
    TransposePlan *p = nullptr;
    for (auto &q : h->transpose_plans)
        if (q.rowptr == rowptr && q.colidx == colidx && q.m == m && q.n == n && q.nnz_hint == nnz_hint && q.base == base) { p = &q; break; }
    if (!p) {
        if (h->transpose_plans.size() >= 8) transpose_drop_plan(h, 0);
        TransposePlan q;
        q.rowptr = rowptr; q.colidx = colidx; q.m = m; q.n = n; q.nnz_hint = nnz_hint; q.base = base;
        RSG_TRY(stored_range(h, m, base, rowptr, &q.first, &q.nnz));
        const size_t cnt = (size_t)std::max(q.nnz, 1);
        if (cudaMalloc(&q.colptr, sizeof(int) * ((size_t)n + 1)) != cudaSuccess || cudaMalloc(&q.rowidx, sizeof(int) * cnt) != cudaSuccess ||
            cudaMalloc(&q.perm, sizeof(int) * cnt) != cudaSuccess || cudaMalloc(&q.values_t, 8 * cnt) != cudaSuccess) {
            (void)cudaGetLastError();
            transpose_free_plan(q);
            return RSG_ALLOC_FAILED;
        }
        const int rc = build_transpose(h, m, n, q.nnz, q.first, base, rowptr, colidx, q.colptr, q.rowidx, q.perm);
        if (rc != RSG_OK) { transpose_free_plan(q); return rc; }
        if (q.nnz > 0) { // longest column of A, from the SpMV plan of the transposed pattern (built here, used by the fallback)
            SpmvPlan *sp = nullptr;
            const int rc2 = spmv_get_plan(h, n, m, q.nnz, base, q.colptr, q.rowidx, false, &sp, (int)sizeof(T));
            q.max_col = rc2 == RSG_OK ? sp->max_row : -1;
        }
        h->transpose_plans.push_back(q);
        p = &h->transpose_plans.back();
    }
    if (p->nnz > 0) {
        if (!values) return RSG_BAD_ARG;
        const int mode = h->opts[RSG_OPT_TRANSPOSE_STATIC];
        if (mode == 0 && p->max_col >= 0 && p->max_col <= kFusedMaxColumn && !h->wait_flags && !h->fuse.ctas) {
            const double avg = (double)p->nnz / std::max(n, 1);
            int lanes = h->opts[RSG_OPT_SPMV_REDUCE];
            if (lanes != 1 && lanes != 2 && lanes != 4 && lanes != 8 && lanes != 16 && lanes != 32)
                lanes = avg <= 8 ? 1 : avg <= 14 ? 2 : avg <= 56 ? 4 : avg <= 160 ? 8 : avg <= 512 ? 16 : 32;
            const int threads = 256;
#define RSG_TF(G)                                                                                                   \
    spmv_t_fused_kernel<T, G, (sizeof(T) == 4 ? 8 : 4)><<<(int)(((long long)n * G + threads - 1) / threads), threads, 0, h->stream>>>(          \
        n, base, p->colptr, p->rowidx, p->perm, values + p->first, x, y, alpha, beta)
            if (lanes == 1) RSG_TF(1); else if (lanes == 2) RSG_TF(2); else if (lanes == 4) RSG_TF(4); else if (lanes == 8) RSG_TF(8); else if (lanes == 16) RSG_TF(16); else RSG_TF(32);
#undef RSG_TF
            h->launches++;
            h->last_spmv_algo = RSG_SPMV_VECTOR;
            h->last_spmv_maxrow = p->max_col;
            RSG_CUDA_TRY(cudaGetLastError());
            return RSG_OK;
        }
        const bool fresh = mode == 1 && p->gathered_from == values && p->gathered_size == (int)sizeof(T);
        if (!fresh) {
            tr_gather_kernel<T><<<std::min((p->nnz + 255) / 256, h->sm_count * 16), 256, 0, h->stream>>>(
                p->nnz, p->perm, values + p->first, static_cast<T *>(p->values_t));
            h->launches++;
            p->gathered_from = values;
            p->gathered_size = (int)sizeof(T);
        }
    }
    // A^T is n x m: its "rows" are A's columns
    return spmv_run<T, T>(h, n, m, p->nnz, base, p->colptr, p->rowidx, static_cast<const T *>(p->values_t), alpha, x, beta, y);
}

template int csr2csc_run<double>(rsg_context *, int, int, int, const int *, const int *, const double *, int *, int *, double *);
template int csr2csc_run<float>(rsg_context *, int, int, int, const int *, const int *, const float *, int *, int *, float *);
template int spmv_transposed_run<double>(rsg_context *, int, int, int, int, const int *, const int *, const double *, double, const double *, double, double *);
template int spmv_transposed_run<float>(rsg_context *, int, int, int, int, const int *, const int *, const float *, float, const float *, float, float *);

} // namespace rsg
