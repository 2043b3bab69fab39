// levels.cu -- "Symbolic" phase: diagonal positions, structural zero-pivot detection and the
// level-set schedules of the lower and upper patterns, all built on the device.  Replaces
// cusparse?csrilu02_analysis and the two csrsv2_analysis calls of the reference harness
// (/root/reference/GPU/ilu0.cu:196-252).
//
// Level of a row = longest dependency chain below it:
//   L: lev[i] = 1 + max{lev[j] : j < i, a_ij stored}   U: same over j > i, from the bottom.
// One thread per row walks its dependencies; a level not yet known is read back as -1 and the
// row simply tries again on the next sweep of its warp (the value is its own ready flag, so no
// fences are needed).  Rows are handed out through an atomic ticket in dependency order, so a
// row only ever waits on rows owned by warps that are already running: no deadlock, no
// grid-wide barrier per level.  The schedule ("level_rows") is a stable radix sort by level.
#include "rsg_internal.cuh"

#include <cub/device/device_radix_sort.cuh>

namespace rsg {

__device__ __forceinline__ int ld_relaxed(const int *p)
{
    int v;
    asm volatile("ld.relaxed.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed(int *p, int v)
{
    asm volatile("st.relaxed.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// diag_pos[i] = zero-based position of a_ii, or -1; stats[0] = min row without diagonal (init INT_MAX)
__global__ void diag_pos_kernel(int n, int base, const int *__restrict__ rowptr, const int *__restrict__ colidx,
                                int *__restrict__ diag_pos, int *stats)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int lo = rowptr[i] - base, hi = rowptr[i + 1] - base;
    atomicMax(stats + 1, hi - lo);
    const int want = i + base;
    int pos = -1;
    // columns are sorted within a row (loader contract): binary search, then confirm
    int a = lo, b = hi;
    while (a < b) {
        const int mid = (a + b) >> 1;
        if (colidx[mid] < want) a = mid + 1; else b = mid;
    }
    if (a < hi && colidx[a] == want) pos = a;
    else { // tolerate unsorted rows: linear scan
        for (int k = lo; k < hi; ++k)
            if (colidx[k] == want) { pos = k; break; }
    }
    diag_pos[i] = pos;
    if (pos < 0) atomicMin(stats, i);
}

// which = 0: lower pattern, rows ascending; which = 1: upper pattern, rows descending.
// ctl[0] = ticket, stats[2] = max level.
__global__ void __launch_bounds__(256)
levels_kernel(int n, int base, int which, const int *__restrict__ rowptr, const int *__restrict__ colidx,
              int *lev, int *ctl, int *stats)
{
    const int lane = threadIdx.x & 31;
    for (;;) {
        int t0 = 0;
        if (lane == 0) t0 = atomicAdd(ctl, 32);
        t0 = __shfl_sync(0xffffffffu, t0, 0);
        if (t0 >= n) break;
        const int t = t0 + lane;
        const bool have = t < n;
        const int i = which == 0 ? t : n - 1 - t;
        int k = 0, ke = 0, mx = 0;
        if (have) { k = rowptr[i] - base; ke = rowptr[i + 1] - base; }
        bool done = !have;
        while (!__all_sync(0xffffffffu, done)) {
            if (!done) {
                while (k < ke) {
                    const int j = colidx[k] - base;
                    const bool dep = which == 0 ? (j < i) : (j > i && j < n);
                    if (dep) {
                        const int l = ld_relaxed(lev + j);
                        if (l < 0) break; // not known yet: try again next sweep
                        mx = max(mx, l + 1);
                    }
                    ++k;
                }
                if (k >= ke) {
                    st_relaxed(lev + i, mx);
                    done = true;
                }
            }
        }
        int wm = have ? mx : 0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) wm = max(wm, __shfl_xor_sync(0xffffffffu, wm, o));
        if (lane == 0) atomicMax(stats + 2, wm);
    }
}

__global__ void iota_kernel(int n, int *a)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) a[i] = i;
}

// level_ptr[l] = first position in the level-sorted key array with key >= l
__global__ void level_ptr_kernel(int n, int nlevels, const int *__restrict__ sorted_lev, int *level_ptr)
{
    const int l = blockIdx.x * blockDim.x + threadIdx.x;
    if (l > nlevels) return;
    int lo = 0, hi = n;
    while (lo < hi) {
        const int mid = lo + ((hi - lo) >> 1);
        if (sorted_lev[mid] < l) lo = mid + 1; else hi = mid;
    }
    level_ptr[l] = lo;
}

int ensure_cub_tmp(rsg_context *h, size_t bytes)
{
    if (bytes <= h->cub_tmp_bytes) return RSG_OK;
    if (h->cub_tmp) cudaFree(h->cub_tmp);
    h->cub_tmp = nullptr; h->cub_tmp_bytes = 0;
    if (cudaMalloc(&h->cub_tmp, bytes) != cudaSuccess) { (void)cudaGetLastError(); return RSG_ALLOC_FAILED; }
    h->cub_tmp_bytes = bytes;
    return RSG_OK;
}

void tri_free_plan(TriPlan &p)
{
    cudaFree(p.diag_pos); cudaFree(p.done); cudaFree(p.ctl);
    for (int w = 0; w < 2; ++w) { cudaFree(p.lev[w]); cudaFree(p.order[w]); cudaFree(p.level_ptr[w]); }
    p = TriPlan();
}

int tri_get_plan(rsg_context *h, int n, int nnz_hint, int base, const int *rowptr, const int *colidx, bool force_rebuild,
                 TriPlan **out)
{
    for (size_t i = 0; i < h->tri_plans.size(); ++i) {
        TriPlan &p = h->tri_plans[i];
        if (p.rowptr == rowptr && p.colidx == colidx && p.n == n && p.nnz_hint == nnz_hint && p.base == base) {
            if (!force_rebuild) { *out = &p; h->last_tri = (int)i; return RSG_OK; }
            tri_free_plan(p);
            h->tri_plans.erase(h->tri_plans.begin() + (long)i);
            break;
        }
    }
    if (h->tri_plans.size() >= 8) {
        tri_free_plan(h->tri_plans.front());
        h->tri_plans.erase(h->tri_plans.begin());
    }
    TriPlan p;
    p.rowptr = rowptr; p.colidx = colidx; p.n = n; p.nnz_hint = nnz_hint; p.base = base;
    cudaStream_t st = h->stream;
    const size_t nn = (size_t)(n > 0 ? n : 1);
    RSG_CUDA_TRY(cudaMalloc(&p.diag_pos, sizeof(int) * nn));
    RSG_CUDA_TRY(cudaMalloc(&p.done, sizeof(int) * nn));
    RSG_CUDA_TRY(cudaMalloc(&p.ctl, sizeof(int) * 8));
    p.bytes = sizeof(int) * (2 * nn + 8);
    RSG_CUDA_TRY(cudaMemsetAsync(p.done, 0, sizeof(int) * nn, st));
    RSG_CUDA_TRY(cudaMemsetAsync(p.ctl, 0, sizeof(int) * 8, st));
    h->pinned_i[0] = 0x7fffffff; h->pinned_i[1] = 0; h->pinned_i[2] = 0; h->pinned_i[3] = 0;
    RSG_CUDA_TRY(cudaMemcpyAsync(h->scratch_i, h->pinned_i, 4 * sizeof(int), cudaMemcpyHostToDevice, st));
    if (n > 0) {
        diag_pos_kernel<<<(n + 255) / 256, 256, 0, st>>>(n, base, rowptr, colidx, p.diag_pos, h->scratch_i);
        h->launches++;
        RSG_CUDA_TRY(cudaMemcpyAsync(h->scratch_i + 4, rowptr, sizeof(int), cudaMemcpyDeviceToDevice, st));
        RSG_CUDA_TRY(cudaMemcpyAsync(h->scratch_i + 5, rowptr + n, sizeof(int), cudaMemcpyDeviceToDevice, st));
    }
    RSG_CUDA_TRY(cudaStreamSynchronize(st)); // the H2D above reads pinned_i: finish before reusing it
    RSG_CUDA_TRY(cudaMemcpyAsync(h->pinned_i, h->scratch_i, 8 * sizeof(int), cudaMemcpyDeviceToHost, st));
    RSG_CUDA_TRY(cudaStreamSynchronize(st));
    p.missing_diag_row = (h->pinned_i[0] == 0x7fffffff) ? -1 : h->pinned_i[0];
    p.max_row = h->pinned_i[1];
    p.nnz = n > 0 ? h->pinned_i[5] - h->pinned_i[4] : 0;
    h->tri_plans.push_back(p);
    h->last_tri = (int)h->tri_plans.size() - 1;
    *out = &h->tri_plans.back();
    return RSG_OK;
}

int tri_build_levels(rsg_context *h, TriPlan *p, int which)
{
    if (which < 0 || which > 1) return RSG_BAD_ARG;
    if (p->have[which]) return RSG_OK;
    const int n = p->n;
    cudaStream_t st = h->stream;
    const size_t nn = (size_t)(n > 0 ? n : 1);
    RSG_CUDA_TRY(cudaMalloc(&p->lev[which], sizeof(int) * nn));
    RSG_CUDA_TRY(cudaMalloc(&p->order[which], sizeof(int) * nn));
    p->bytes += 2 * sizeof(int) * nn;
    if (n == 0) {
        p->nlevels[which] = 0;
        RSG_CUDA_TRY(cudaMalloc(&p->level_ptr[which], sizeof(int)));
        RSG_CUDA_TRY(cudaMemsetAsync(p->level_ptr[which], 0, sizeof(int), st));
        p->level_ptr_host[which].assign(1, 0);
        p->have[which] = true;
        return RSG_OK;
    }
    RSG_CUDA_TRY(cudaMemsetAsync(p->lev[which], 0xff, sizeof(int) * nn, st)); // -1 = unknown
    RSG_CUDA_TRY(cudaMemsetAsync(p->ctl, 0, sizeof(int) * 8, st));
    RSG_CUDA_TRY(cudaMemsetAsync(h->scratch_i, 0, sizeof(int) * 4, st));
    {
        int per_sm = 0;
        RSG_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, levels_kernel, 256, 0));
        const int grid = std::max(1, std::min(h->sm_count * std::max(per_sm, 1), (n + 255) / 256));
        levels_kernel<<<grid, 256, 0, st>>>(n, p->base, which, p->rowptr, p->colidx, p->lev[which], p->ctl, h->scratch_i);
        h->launches++;
    }
    RSG_CUDA_TRY(cudaMemcpyAsync(h->pinned_i, h->scratch_i, 4 * sizeof(int), cudaMemcpyDeviceToHost, st));
    RSG_CUDA_TRY(cudaStreamSynchronize(st));
    const int nlev = h->pinned_i[2] + 1;
    p->nlevels[which] = nlev;
    // stable sort of rows by level => rows ascending inside each level
    int *keys_out = nullptr, *vals_in = nullptr;
    RSG_CUDA_TRY(cudaMalloc(&keys_out, sizeof(int) * nn));
    RSG_CUDA_TRY(cudaMalloc(&vals_in, sizeof(int) * nn));
    iota_kernel<<<(n + 255) / 256, 256, 0, st>>>(n, vals_in);
    h->launches++;
    int end_bit = 1;
    while (end_bit < 31 && (1 << end_bit) < nlev) ++end_bit;
    size_t tmp = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, tmp, p->lev[which], keys_out, vals_in, p->order[which], n, 0, end_bit, st);
    int rc = ensure_cub_tmp(h, tmp);
    if (rc == RSG_OK) {
        cudaError_t e = cub::DeviceRadixSort::SortPairs(h->cub_tmp, tmp, p->lev[which], keys_out, vals_in,
                                                         p->order[which], n, 0, end_bit, st);
        if (e != cudaSuccess) rc = RSG_CUDA_ERROR;
        h->launches += 3;
    }
    if (rc == RSG_OK) {
        if (cudaMalloc(&p->level_ptr[which], sizeof(int) * ((size_t)nlev + 1)) != cudaSuccess) rc = RSG_ALLOC_FAILED;
    }
    if (rc == RSG_OK) {
        p->bytes += sizeof(int) * ((size_t)nlev + 1);
        level_ptr_kernel<<<(nlev + 1 + 255) / 256, 256, 0, st>>>(n, nlev, keys_out, p->level_ptr[which]);
        h->launches++;
        p->level_ptr_host[which].resize((size_t)nlev + 1);
        if (cudaMemcpyAsync(p->level_ptr_host[which].data(), p->level_ptr[which], sizeof(int) * ((size_t)nlev + 1),
                            cudaMemcpyDeviceToHost, st) != cudaSuccess) rc = RSG_CUDA_ERROR;
    }
    if (cudaStreamSynchronize(st) != cudaSuccess) rc = RSG_CUDA_ERROR;
    cudaFree(keys_out); cudaFree(vals_in);
    if (rc != RSG_OK) { (void)cudaGetLastError(); return rc; }
    p->have[which] = true;
    return RSG_OK;
}

} // namespace rsg
