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

// diag_pos[i] = zero-based position of a_ii, or -1; stats[0] = min row without diagonal (init INT_MAX);
// stats[2] = 1 when some row's columns are not ascending (the factorization and the solves split a row at its
// diagonal and binary-search it: such input is refused with RSG_BAD_ARG instead of giving wrong factors)
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
    bool descending = false;
    for (int k = lo + 1; k < hi; ++k) descending |= colidx[k] < colidx[k - 1];
    if (descending) stats[2] = 1;
}

// Both patterns in ONE launch: CTAs with an even index walk the lower pattern (rows ascending),
// odd ones the upper pattern (rows descending); `only` = 0/1 restricts the launch to one pattern,
// -1 runs both side by side (the two dependency chains overlap instead of running back to back).
// A thread keeps a batch of B dependencies (their columns) in registers and fetches the B levels
// in one round trip; a batch with an unknown level is simply fetched again on the next sweep.
// ctl[0] / ctl[4] = tickets of L / U, stats[2] / stats[3] = max level of L / U.
template <int B>
__global__ void __launch_bounds__(256)
levels_kernel(int n, int base, int only, const int *__restrict__ rowptr, const int *__restrict__ colidx,
              int *lev_l, int *lev_u, int *crit_l, int *crit_u, int *ctl, int *stats)
{
    const int lane = threadIdx.x & 31;
    const int which = only >= 0 ? only : (int)(blockIdx.x & 1);
    int *lev = which == 0 ? lev_l : lev_u;
    int *crit = which == 0 ? crit_l : crit_u;
    int *ticket = ctl + 4 * which;
    int wmax = 0;
    for (;;) {
        int t0 = 0;
        if (lane == 0) t0 = atomicAdd(ticket, 32);
        t0 = __shfl_sync(0xffffffffu, t0, 0);
        if (t0 >= n) break;
        const int t = t0 + lane;
        const bool have = t < n;
        const int i = which == 0 ? t : n - 1 - t;
        int k = 0, ke = 0, mx = 0, mxcol = -1; // mxcol: the dependency that finishes last (deepest level)
        if (have) {
            // only the dependency part of the (column-sorted) row is walked
            k = rowptr[i] - base;
            ke = rowptr[i + 1] - base;
        }
        int col[B];
        int nb = 0; // dependencies of the current batch (columns in col[0..nb))
        bool done = !have;
        while (!__all_sync(0xffffffffu, done)) {
            if (!done) {
                for (;;) {
                    if (nb == 0) {
                        // next batch: skip entries that are not dependencies
                        while (k < ke && nb < B) {
                            const int j = colidx[k] - base;
                            const bool dep = which == 0 ? (j < i) : (j > i && j < n);
                            if (dep) {
#pragma unroll
                                for (int u = 0; u < B; ++u)
                                    if (u == nb) col[u] = j;
                                ++nb;
                            }
                            ++k;
                        }
                        if (nb == 0) { crit[i] = mxcol; st_relaxed(lev + i, mx); done = true; break; }
                    }
                    int l[B];
#pragma unroll
                    for (int u = 0; u < B; ++u)
                        if (u < nb) l[u] = ld_relaxed(lev + col[u]);
                    int lo = 0, m2 = mx, c2 = mxcol;
#pragma unroll
                    for (int u = 0; u < B; ++u)
                        if (u < nb) {
                            lo = min(lo, l[u]);
                            if (l[u] + 1 > m2) { m2 = l[u] + 1; c2 = col[u]; }
                        }
                    if (lo < 0) break; // a level is not known yet: fetch the batch again next sweep
                    mx = m2;
                    mxcol = c2;
                    nb = 0;
                }
            }
        }
        if (have) wmax = max(wmax, mx);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) wmax = max(wmax, __shfl_xor_sync(0xffffffffu, wmax, o));
    if (lane == 0 && wmax > 0) atomicMax(stats + 2 + which, wmax);
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

void tri_free_chain(TriPlan &p)
{
    for (int w = 0; w < 2; ++w) {
        if (p.tmeta[w]) cudaFree(p.tmeta[w]);
        if (p.chunks[w]) cudaFree(p.chunks[w]);
        p.tmeta[w] = nullptr; p.chunks[w] = nullptr; p.nchunks[w] = 0;
    }
}

void tri_free_plan(TriPlan &p)
{
    tri_free_chain(p);
    if (p.slab) cudaFree(p.slab);
    p = TriPlan();
}

// Every per-plan array lives in one slab (a cudaMalloc / cudaFree pair costs more than the whole
// level computation, and the reference harness re-runs the analysis for every factorization).
static int tri_carve(TriPlan &p, int n)
{
    const size_t nn = ((size_t)(n > 0 ? n : 1) + 31) & ~(size_t)31; // 128-byte aligned sub-arrays
    const size_t ints = 12 * nn + 64;
    if (!p.slab || p.slab_n != n) {
        if (p.slab) cudaFree(p.slab);
        p.slab = nullptr;
        if (cudaMalloc(&p.slab, sizeof(int) * ints) != cudaSuccess) { (void)cudaGetLastError(); return RSG_ALLOC_FAILED; }
        p.slab_n = n;
        p.epoch = 0;
    }
    int *q = p.slab;
    p.ctl = q; q += 32;
    p.diag_pos = q; q += nn;
    p.done = q; q += nn;
    p.lev[0] = q; q += nn;
    p.lev[1] = q; q += nn; // contiguous with lev[0]: one memset initialises both
    p.order[0] = q; q += nn;
    p.order[1] = q; q += nn;
    p.level_ptr[0] = q; q += nn + 16; // at most n+1 entries each
    p.level_ptr[1] = q; q += nn + 16;
    p.crit[0] = q; q += nn;
    p.crit[1] = q; q += nn;
    p.sort_keys = q; q += nn;
    p.sort_vals = q;
    p.bytes = sizeof(int) * ints;
    return RSG_OK;
}

int tri_get_plan(rsg_context *h, int n, int nnz_hint, int base, const int *rowptr, const int *colidx, bool force_rebuild,
                 TriPlan **out)
{
    TriPlan p;
    for (size_t i = 0; i < h->tri_plans.size(); ++i) {
        TriPlan &q = h->tri_plans[i];
        if (q.rowptr == rowptr && q.colidx == colidx && q.n == n && q.nnz_hint == nnz_hint && q.base == base) {
            if (!force_rebuild) { *out = &q; h->last_tri = (int)i; return RSG_OK; }
            // re-analysis of the same arrays: keep the slab (and its epoch), rebuild the contents
            p.slab = q.slab; p.slab_n = q.slab_n; p.epoch = q.epoch;
            tri_free_chain(q);
            h->tri_plans.erase(h->tri_plans.begin() + (long)i);
            break;
        }
    }
    if (h->tri_plans.size() >= 8) {
        tri_free_plan(h->tri_plans.front());
        h->tri_plans.erase(h->tri_plans.begin());
    }
    p.rowptr = rowptr; p.colidx = colidx; p.n = n; p.nnz_hint = nnz_hint; p.base = base;
    cudaStream_t st = h->stream;
    const bool fresh = (p.slab == nullptr || p.slab_n != n);
    RSG_TRY(tri_carve(p, n));
    if (fresh) RSG_CUDA_TRY(cudaMemsetAsync(p.done, 0, sizeof(int) * (size_t)(n > 0 ? n : 1), st));
    RSG_CUDA_TRY(cudaMemsetAsync(p.ctl, 0, sizeof(int) * 32, st));
    // stats: [0] min row without diagonal, [1] longest row, [4],[5] first / last row pointer
    h->pinned_i[0] = 0x7fffffff; h->pinned_i[1] = 0; h->pinned_i[2] = 0; h->pinned_i[3] = 0;
    h->pinned_i[4] = 0; h->pinned_i[5] = 0;
    RSG_CUDA_TRY(cudaMemcpyAsync(h->scratch_i, h->pinned_i, 8 * sizeof(int), cudaMemcpyHostToDevice, st));
    if (n > 0) {
        diag_pos_kernel<<<(n + 255) / 256, 256, 0, st>>>(n, base, rowptr, colidx, p.diag_pos, h->scratch_i);
        h->launches++;
        RSG_CUDA_TRY(cudaMemcpyAsync(h->scratch_i + 4, rowptr, sizeof(int), cudaMemcpyDeviceToDevice, st));
        RSG_CUDA_TRY(cudaMemcpyAsync(h->scratch_i + 5, rowptr + n, sizeof(int), cudaMemcpyDeviceToDevice, st));
    }
    RSG_CUDA_TRY(cudaMemcpyAsync(h->pinned_i + 16, h->scratch_i, 8 * sizeof(int), cudaMemcpyDeviceToHost, st));
    RSG_CUDA_TRY(cudaStreamSynchronize(st));
    const int *r = h->pinned_i + 16;
    p.missing_diag_row = (r[0] == 0x7fffffff) ? -1 : r[0];
    p.max_row = r[1];
    p.unsorted = r[2] != 0;
    p.nnz = n > 0 ? r[5] - r[4] : 0;
    h->tri_plans.push_back(p);
    h->last_tri = (int)h->tri_plans.size() - 1;
    *out = &h->tri_plans.back();
    return RSG_OK;
}

// which = 0 (L), 1 (U) or -1 (both patterns, one launch of the level kernel).
int tri_build_levels(rsg_context *h, TriPlan *p, int which)
{
    if (which < -1 || which > 1) return RSG_BAD_ARG;
    bool need[2] = {(which == 0 || which == -1) && !p->have[0], (which == 1 || which == -1) && !p->have[1]};
    if (!need[0] && !need[1]) return RSG_OK;
    const int n = p->n;
    cudaStream_t st = h->stream;
    if (n == 0) {
        for (int w = 0; w < 2; ++w)
            if (need[w]) {
                p->nlevels[w] = 0;
                RSG_CUDA_TRY(cudaMemsetAsync(p->level_ptr[w], 0, sizeof(int), st));
                p->level_ptr_host[w].assign(1, 0);
                p->have[w] = true;
            }
        return RSG_OK;
    }
    const int only = (need[0] && need[1]) ? -1 : (need[0] ? 0 : 1);
    for (int w = 0; w < 2; ++w)
        if (need[w]) RSG_CUDA_TRY(cudaMemsetAsync(p->lev[w], 0xff, sizeof(int) * (size_t)n, st)); // -1 = unknown
    RSG_CUDA_TRY(cudaMemsetAsync(p->ctl, 0, sizeof(int) * 8, st));
    RSG_CUDA_TRY(cudaMemsetAsync(h->scratch_i, 0, sizeof(int) * 4, st));
    {
        constexpr int B = 8;
        int per_sm = 0;
        RSG_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, levels_kernel<B>, 256, 0));
        const int ctas_per_pattern = std::max(1, std::min(h->sm_count * std::max(per_sm, 1) / (only < 0 ? 2 : 1), (n + 255) / 256));
        const int grid = only < 0 ? 2 * ctas_per_pattern : ctas_per_pattern;
        levels_kernel<B><<<grid, 256, 0, st>>>(n, p->base, only, p->rowptr, p->colidx, p->lev[0], p->lev[1], p->crit[0], p->crit[1], p->ctl,
                                                   h->scratch_i);
        h->launches++;
    }
    RSG_CUDA_TRY(cudaMemcpyAsync(h->pinned_i + 24, h->scratch_i, 4 * sizeof(int), cudaMemcpyDeviceToHost, st));
    iota_kernel<<<(n + 255) / 256, 256, 0, st>>>(n, p->sort_vals);
    h->launches++;
    RSG_CUDA_TRY(cudaStreamSynchronize(st));
    int rc = RSG_OK;
    for (int w = 0; w < 2 && rc == RSG_OK; ++w) {
        if (!need[w]) continue;
        const int nlev = h->pinned_i[24 + 2 + w] + 1;
        p->nlevels[w] = nlev;
        // stable sort of rows by level => rows ascending inside each level
        int end_bit = 1;
        while (end_bit < 31 && (1 << end_bit) < nlev) ++end_bit;
        size_t tmp = 0;
        cub::DeviceRadixSort::SortPairs(nullptr, tmp, p->lev[w], p->sort_keys, p->sort_vals, p->order[w], n, 0, end_bit, st);
        rc = ensure_cub_tmp(h, tmp);
        if (rc != RSG_OK) break;
        if (cub::DeviceRadixSort::SortPairs(h->cub_tmp, tmp, p->lev[w], p->sort_keys, p->sort_vals, p->order[w], n, 0, end_bit,
                                            st) != cudaSuccess) { rc = RSG_CUDA_ERROR; break; }
        h->launches += 3;
        level_ptr_kernel<<<(nlev + 1 + 255) / 256, 256, 0, st>>>(n, nlev, p->sort_keys, p->level_ptr[w]);
        h->launches++;
        p->level_ptr_host[w].resize((size_t)nlev + 1);
        if (cudaMemcpyAsync(p->level_ptr_host[w].data(), p->level_ptr[w], sizeof(int) * ((size_t)nlev + 1),
                            cudaMemcpyDeviceToHost, st) != cudaSuccess) rc = RSG_CUDA_ERROR;
    }
    if (cudaStreamSynchronize(st) != cudaSuccess && rc == RSG_OK) rc = RSG_CUDA_ERROR;
    if (rc != RSG_OK) { (void)cudaGetLastError(); return rc; }
    for (int w = 0; w < 2; ++w)
        if (need[w]) p->have[w] = true;
    return RSG_OK;
}

} // namespace rsg
