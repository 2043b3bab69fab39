// spmv.cu -- CSR SpMV  y = alpha*A*x + beta*y  for sm_100a (B200), replacing the cusparseSpMV
// call group of the reference harness (/root/reference/GPU/spmv.cu:122-186).
//
// Kernels (no tensor cores: 2 flop per 12 B, the bound is HBM):
//   rowblock  The matrix is cut on device into CTA tiles of ~TILE merged items (rows + entries)
//             rounded to row boundaries.  A CTA streams its entries with aligned 128-bit
//             loads indexed on the GLOBAL entry index (row starts are arbitrary, the stream is
//             not), multiplies by the gathered x (read-only path, L1/L2 resident) and parks the
//             products in shared memory; rows are then reduced out of shared memory by a
//             per-tile row-length bin: thread-per-row (short rows), 4/8/16/32-lane
//             sub-warp-per-row with shuffle reduction (medium rows).  A row that does not fit
//             the tile is reduced CTA-per-row straight from global memory.
//   merge     Merge-path (rows+entries split evenly over CTAs and threads, carry fix-up):
//             for skewed row lengths, where one dense row would pin a single CTA.
//   window / tile   the TMA-staged kernels (spmv_window.cu, spmv_tile.cu); this file holds the plan logic that picks them
//   vector    plain sub-warp-per-row, batched scalar loads, any alignment (scattered columns, small matrices)
//   scalar    plain thread-per-row (fallback / cross-check)
// TX is the vector / accumulation type, TV the stored-value type: (double,double) fp64,
// (float,float) fp32, (float,double) the mixed variant.
#include "spmv_device.cuh"

#include <cub/block/block_scan.cuh>

namespace rsg {

// ------------------------------------------------------------------ row-block kernel
template <int TILE> struct RowBlockShape {
    static constexpr int CAP = TILE + TILE / 8; // staged entries per pass
};

template <typename TV, typename TX, int THREADS, int TILE, int HINT, bool VEC>
__global__ void __launch_bounds__(THREADS)
spmv_rowblock_kernel(int base, const int *__restrict__ rowptr, const int *__restrict__ colidx,
                     const TV *__restrict__ values, const TX *__restrict__ x, TX *__restrict__ y,
                     TX alpha, TX beta, const int *__restrict__ blk_row, int reduce_opt)
{
    constexpr int CAP = RowBlockShape<TILE>::CAP;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    TX *s_prod = reinterpret_cast<TX *>(smem_raw);            // CAP + 4
    int *s_rp = reinterpret_cast<int *>(s_prod + CAP + 4);    // TILE + 1
    const int b = blockIdx.x;
    const int r0 = blk_row[b], r1 = blk_row[b + 1];
    const int nr = r1 - r0;
    if (nr <= 0) return;
    uint64_t pol = 0;
    if (HINT == 2) pol = make_evict_first_policy();
    for (int i = threadIdx.x; i <= nr; i += THREADS) s_rp[i] = rowptr[r0 + i] - base;
    __syncthreads();
    int cur = 0;
    while (cur < nr) {
        const int k0 = s_rp[cur];
        const int sb = k0 & ~3;
        int pe;
        if (s_rp[nr] - sb <= CAP) {
            pe = nr;
        } else { // largest pe in [cur,nr] with s_rp[pe]-sb <= CAP
            int lo = cur, hi = nr;
            while (lo < hi) {
                const int mid = (lo + hi + 1) >> 1;
                if (s_rp[mid] - sb <= CAP) lo = mid; else hi = mid - 1;
            }
            pe = lo;
        }
        if (pe == cur) { // this row alone exceeds the staging capacity: CTA-per-row
            const TX tot = cta_row_sum<TV, TX, THREADS, HINT, VEC>(k0, s_rp[cur + 1], colidx, values, x, base, s_prod, pol);
            if (threadIdx.x == 0) store_y(y, r0 + cur, tot, alpha, beta);
            cur += 1;
            continue;
        }
        const int k1 = s_rp[pe];
        stage_products<TV, TX, THREADS, HINT, VEC>(k0, k1, colidx, values, x, base, s_prod, pol);
        __syncthreads();
        const int nrows = pe - cur;
        int g = reduce_opt;
        if (g == 0) { // row-length bin of this tile
            const int avg = (k1 - k0) / nrows;
            g = avg <= 12 ? 1 : avg <= 48 ? 4 : avg <= 96 ? 8 : avg <= 256 ? 16 : 32;
        }
        const int *rp = s_rp + cur;
        switch (g) {
        case 1: reduce_rows<TX, THREADS, 1>(s_prod, rp, sb, nrows, r0 + cur, y, alpha, beta); break;
        case 2: reduce_rows<TX, THREADS, 2>(s_prod, rp, sb, nrows, r0 + cur, y, alpha, beta); break;
        case 4: reduce_rows<TX, THREADS, 4>(s_prod, rp, sb, nrows, r0 + cur, y, alpha, beta); break;
        case 8: reduce_rows<TX, THREADS, 8>(s_prod, rp, sb, nrows, r0 + cur, y, alpha, beta); break;
        case 16: reduce_rows<TX, THREADS, 16>(s_prod, rp, sb, nrows, r0 + cur, y, alpha, beta); break;
        default: reduce_rows<TX, THREADS, 32>(s_prod, rp, sb, nrows, r0 + cur, y, alpha, beta); break;
        }
        cur = pe;
        if (cur < nr) __syncthreads();
    }
}


// first row r in [0,m] with (rowptr[r]-rowptr[0]) + r >= target  (merged-item coordinate)
__device__ __forceinline__ int first_row_at_item(const int *__restrict__ rowptr, int m, long long target)
{
    const int rp0 = rowptr[0];
    int lo = 0, hi = m;
    while (lo < hi) {
        const int mid = lo + ((hi - lo) >> 1);
        if ((long long)(rowptr[mid] - rp0) + mid < target) lo = mid + 1; else hi = mid;
    }
    return lo;
}

__global__ void rowblock_partition_kernel(int m, const int *__restrict__ rowptr, int tile, int nblk, int *blk_row)
{
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b > nblk) return;
    blk_row[b] = (b == nblk) ? m : first_row_at_item(rowptr, m, (long long)b * tile);
}

// ------------------------------------------------------------------ merge-path kernel
template <typename TX> struct CarryPair {
    int f; // 1: a row ended inside this thread's span (v restarts there)
    TX v;
};
template <typename TX> struct CarryOp {
    __device__ __forceinline__ CarryPair<TX> operator()(const CarryPair<TX> &a, const CarryPair<TX> &b) const
    {
        CarryPair<TX> r;
        r.f = a.f | b.f;
        r.v = b.f ? b.v : a.v + b.v;
        return r;
    }
};

// coords[t] = (rows consumed, absolute entry index) where diagonal t*items crosses the merge of
// row-end offsets A[i] = rowptr[i+1]-base and entry indices B[j] = first+j.
__global__ void merge_coords_kernel(int m, int nnz, int base, const int *__restrict__ rowptr, int items,
                                    int ntiles, int2 *coords)
{
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t > ntiles) return;
    const int first = rowptr[0] - base;
    long long d = (long long)t * items;
    const long long total = (long long)m + nnz;
    if (d > total) d = total;
    int lo = (int)(d > nnz ? d - nnz : 0), hi = (int)(d < m ? d : m);
    while (lo < hi) {
        const int mid = lo + ((hi - lo) >> 1);
        if ((long long)(rowptr[mid + 1] - base) <= (long long)first + (d - 1 - mid)) lo = mid + 1; else hi = mid;
    }
    coords[t] = make_int2(lo, first + (int)(d - lo));
}

template <typename TV, typename TX, int THREADS, int IPT, int HINT, bool VEC>
__global__ void __launch_bounds__(THREADS)
spmv_merge_kernel(int m, int base, const int *__restrict__ rowptr, const int *__restrict__ colidx,
                  const TV *__restrict__ values, const TX *__restrict__ x, TX *__restrict__ y, TX alpha, TX beta,
                  const int2 *__restrict__ coords, int *__restrict__ carry_row, TX *__restrict__ carry_val)
{
    constexpr int ITEMS = THREADS * IPT;
    typedef cub::BlockScan<CarryPair<TX>, THREADS> Scan;
    __shared__ __align__(16) TX s_prod[ITEMS + 4];
    __shared__ int s_end[ITEMS + 1];
    __shared__ typename Scan::TempStorage s_scan;
    const int tid = threadIdx.x, b = blockIdx.x;
    const int2 c0 = coords[b], c1 = coords[b + 1];
    const int row0 = c0.x, nz0 = c0.y, nr = c1.x - c0.x, nn = c1.y - c0.y;
    uint64_t pol = 0;
    if (HINT == 2) pol = make_evict_first_policy();
    for (int i = tid; i < nr; i += THREADS) s_end[i] = rowptr[row0 + 1 + i] - base;
    stage_products<TV, TX, THREADS, HINT, VEC>(nz0, nz0 + nn, colidx, values, x, base, s_prod, pol);
    __syncthreads();
    const TX *prod = s_prod + (nz0 & 3); // entry nz0+j sits at s_prod[(nz0&3)+j]
    // this thread's start on the tile-local merge path
    int d = tid * IPT;
    const int total = nr + nn;
    if (d > total) d = total;
    int lo = d > nn ? d - nn : 0, hi = d < nr ? d : nr;
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (s_end[mid] <= nz0 + (d - 1 - mid)) lo = mid + 1; else hi = mid;
    }
    int i = lo, j = d - lo;
    TX sum = 0, head_val = 0;
    int head_row = -1;
    const int steps = min(IPT, total - d);
    for (int it = 0; it < steps; ++it) {
        if (j < nn && (i >= nr || s_end[i] > nz0 + j)) {
            sum += prod[j];
            ++j;
        } else { // row row0+i ends here
            if (head_row < 0) { head_row = row0 + i; head_val = sum; }
            else store_y(y, row0 + i, sum, alpha, beta);
            sum = 0;
            ++i;
        }
    }
    CarryPair<TX> in, ex, agg, init;
    in.f = head_row >= 0; in.v = sum;
    init.f = 0; init.v = 0;
    Scan(s_scan).ExclusiveScan(in, ex, init, CarryOp<TX>(), agg);
    if (head_row >= 0) store_y(y, head_row, ex.v + head_val, alpha, beta);
    if (tid == 0) { // the row still open at the end of the tile is row0+nr
        const int open = row0 + nr;
        carry_row[b] = open < m ? open : -1;
        carry_val[b] = agg.v;
    }
}

// adds the carried partial sums: one thread per run of equal carry_row, fixed order
template <typename TX>
__global__ void merge_fixup_kernel(int ntiles, const int *__restrict__ carry_row, const TX *__restrict__ carry_val,
                                   TX *__restrict__ y, TX alpha)
{
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= ntiles) return;
    const int r = carry_row[t];
    if (r < 0 || (t > 0 && carry_row[t - 1] == r)) return;
    TX s = carry_val[t];
    for (int u = t + 1; u < ntiles && carry_row[u] == r; ++u) s += carry_val[u];
    y[r] += alpha * s;
}

// ------------------------------------------------------------------ plain kernels
template <typename TV, typename TX>
__global__ void spmv_scalar_kernel(int m, int base, const int *__restrict__ rowptr, const int *__restrict__ colidx,
                                   const TV *__restrict__ values, const TX *__restrict__ x, TX *__restrict__ y,
                                   TX alpha, TX beta)
{
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= m) return;
    TX s = 0;
    for (int k = rowptr[r] - base; k < rowptr[r + 1] - base; ++k) s += (TX)values[k] * __ldg(x + (colidx[k] - base));
    store_y(y, r, s, alpha, beta);
}

// G lanes per row straight from global memory.  Entries are taken in batches of four per lane (all index loads, then
// all value / x loads, then the multiply-adds), the tail as one predicated batch, so a lane keeps eight gathers in
// flight instead of two; summation order is fixed (lane-strided partial sums, xor tree).
template <typename TV, typename TX, int G, int B>
__global__ void __launch_bounds__(256)
spmv_vector_kernel(int m, int base, const int *__restrict__ rowptr, const int *__restrict__ colidx,
                   const TV *__restrict__ values, const TX *__restrict__ x, TX *__restrict__ y, TX alpha, TX beta)
{
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    const int r = t / G, lane = t % G;
    TX s = 0;
    if (r < m) {
        const int e = rowptr[r + 1] - base;
        int k = rowptr[r] - base + lane;
        for (; k + (B - 1) * G < e; k += B * G) {
            int cc[B];
            TX aa[B], xx[B];
#pragma unroll
            for (int j = 0; j < B; ++j) { cc[j] = colidx[k + j * G] - base; aa[j] = (TX)values[k + j * G]; }
#pragma unroll
            for (int j = 0; j < B; ++j) xx[j] = __ldg(x + cc[j]);
#pragma unroll
            for (int j = 0; j < B; ++j) s += aa[j] * xx[j];
        }
        if (k < e) {
            int cc[B];
            TX aa[B], xx[B];
#pragma unroll
            for (int j = 0; j < B; ++j) {
                const bool ok = k + j * G < e;
                cc[j] = ok ? colidx[k + j * G] - base : 0;
                aa[j] = ok ? (TX)values[k + j * G] : (TX)0;
            }
#pragma unroll
            for (int j = 0; j < B; ++j) xx[j] = (k + j * G < e) ? __ldg(x + cc[j]) : (TX)0;
#pragma unroll
            for (int j = 0; j < B; ++j)
                if (k + j * G < e) s += aa[j] * xx[j];
        }
    }
#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (r < m && lane == 0) store_y(y, r, s, alpha, beta);
}

// lanes per row of the VECTOR kernel by average row length (measured: profiles/r1_vector_lanes_sweep.txt -- fewer
// lanes than entries/4 would suggest: what counts is gathers in flight per lane, and each lane batches four entries)
static inline int vector_lanes(double avg) { return avg <= 15.5 ? 2 : avg <= 34 ? 4 : avg <= 128 ? 8 : avg <= 512 ? 16 : 32; }

// ------------------------------------------------------------------ statistics
// out[0] = max row length, out[1] = rowptr[0], out[2] = rowptr[m]
__global__ void row_stats_kernel(int m, const int *__restrict__ rowptr, int *out)
{
    int mx = 0;
    for (int r = blockIdx.x * blockDim.x + threadIdx.x; r < m; r += gridDim.x * blockDim.x)
        mx = max(mx, rowptr[r + 1] - rowptr[r]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if ((threadIdx.x & 31) == 0 && mx > 0) atomicMax(out, mx);
    if (blockIdx.x == 0 && threadIdx.x == 0) { out[1] = rowptr[0]; out[2] = rowptr[m]; }
}

// How scattered are the x gathers of NEIGHBOURING rows?  One warp per sampled group of 32 consecutive rows: lane l
// takes the middle entry of row r0+l; out[3] += distinct 128-byte lines (of 8-byte x) among them, out[4] += rows
// looked at.  A stencil or FEM ordering gives ~0.1 lines per row (thread-per-row gathers coalesce), hashed columns ~1.
__global__ void scatter_stats_kernel(int m, int base, int groups, const int *__restrict__ rowptr, const int *__restrict__ colidx, int *out)
{
    const int g = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (g >= groups) return;
    const int ngroups32 = (m + 31) / 32;
    const int r = (int)(((long long)g * ngroups32) / groups) * 32 + lane;
    int line = -1 - lane; // distinct dummies for absent rows
    bool valid = false;
    if (r < m) {
        const int a = rowptr[r], b = rowptr[r + 1];
        if (b > a) { line = (colidx[a - base + ((b - a) >> 1)] - base) >> 4; valid = true; }
    }
    const unsigned peers = __match_any_sync(0xffffffffu, line);
    const bool leader = valid && (lane == __ffs(peers) - 1);
    const int distinct = __popc(__ballot_sync(0xffffffffu, leader)), rows = __popc(__ballot_sync(0xffffffffu, valid));
    if (lane == 0 && rows > 0) { atomicAdd(out + 3, distinct); atomicAdd(out + 4, rows); }
}

__global__ void row_partition_kernel(int m, int base, const int *__restrict__ rowptr, int ngpu, int balanced, int *starts)
{
    const int g = threadIdx.x;
    if (g > ngpu) return;
    if (g == 0) { starts[0] = 0; return; }
    if (g == ngpu) { starts[g] = m; return; }
    if (!balanced) { starts[g] = (int)(((long long)g * m + ngpu - 1) / ngpu); return; }
    const long long nnz = (long long)rowptr[m] - rowptr[0];
    const long long target = (long long)g * nnz / ngpu + (rowptr[0] - base);
    int lo = 0, hi = m;
    while (lo < hi) {
        const int mid = lo + ((hi - lo) >> 1);
        if ((long long)rowptr[mid] - base < target) lo = mid + 1; else hi = mid;
    }
    starts[g] = lo;
}

// ------------------------------------------------------------------ host side
struct Variant { int threads, tile; };
static const Variant kVariants[] = {{256, 4096}, {128, 2048}, {256, 2048}, {512, 8192}, {128, 1024}};
static const int kNumVariants = (int)(sizeof(kVariants) / sizeof(kVariants[0]));
static const int kMergeThreads = 256, kMergeIpt = 8;

static long long g_plans_freed = 0; // bumps whenever device arrays of a plan go away (captured graphs key on it)
long long spmv_plans_freed() { return g_plans_freed; }

void spmv_free_plan(SpmvPlan &p)
{
    ++g_plans_freed;
    cudaFree(p.blk_row); cudaFree(p.wdesc); cudaFree(p.widx); cudaFree(p.pdesc); cudaFree(p.rp0); cudaFree(p.seg); cudaFree(p.lrow); cudaFree(p.seg_part); cudaFree(p.seg_cnt); cudaFree(p.coords); cudaFree(p.carry_row); cudaFree(p.carry_val);
    p = SpmvPlan();
}

static int plan_signature(const rsg_context *h)
{
    int sig = h->opts[RSG_OPT_SPMV_ALGO] * 1000003 + h->opts[RSG_OPT_SPMV_VARIANT] * 10007 + h->opts[RSG_OPT_SPMV_LONG_ROW] +
              h->opts[RSG_OPT_SPMV_ROWS_PER_THREAD] * 7919 + h->opts[RSG_OPT_SPMV_REDUCE] * 104729 +
              h->opts[RSG_OPT_SPMV_HALO_START] * 31 + (h->opts[RSG_OPT_SPMV_X_HINT] == 4 ? 1299709 : 0);
    sig = sig * 31 + h->opts[RSG_OPT_SPMV_WIN_BYTES];
    sig = sig * 31 + h->opts[RSG_OPT_SPMV_WIN_STREAM_PCT];
    sig = sig * 31 + h->opts[RSG_OPT_SPMV_WIN_GAP];
    sig = sig * 31 + (int)(((uintptr_t)h->plan_send_pos) >> 4); // a plan for chained products marks its producer tiles
    return sig;
}

int spmv_get_plan(rsg_context *h, int m, int n, int nnz_hint, int base, const int *rowptr, const int *colidx,
                  bool force_rebuild, SpmvPlan **out, int pkey)
{
    const int vbytes = pkey & 15, xbytes = (pkey >> 4) ? (pkey >> 4) : (pkey & 15);
    const int sig = plan_signature(h) + (vbytes + 16 * xbytes) * 15485863; // tile shapes depend on the element sizes
    for (size_t i = 0; i < h->spmv_plans.size(); ++i) {
        SpmvPlan &p = h->spmv_plans[i];
        if (p.rowptr == rowptr && p.colidx == colidx && p.m == m && p.n == n && p.nnz_hint == nnz_hint && p.base == base &&
            p.opt_sig == sig) {
            if (!force_rebuild) { *out = &p; return RSG_OK; }
            spmv_free_plan(p);
            h->spmv_plans.erase(h->spmv_plans.begin() + (long)i);
            break;
        }
    }
    if (h->spmv_plans.size() >= 64) { // bounded cache: drop the oldest
        spmv_free_plan(h->spmv_plans.front());
        h->spmv_plans.erase(h->spmv_plans.begin());
    }
    SpmvPlan p;
    p.rowptr = rowptr; p.colidx = colidx; p.m = m; p.n = n; p.nnz_hint = nnz_hint; p.base = base; p.opt_sig = sig;
    p.vbytes = vbytes; p.xbytes = xbytes;
    cudaStream_t st = h->stream;
    RSG_CUDA_TRY(cudaMemsetAsync(h->scratch_i, 0, 8 * sizeof(int), st));
    if (m > 0) {
        const int blocks = std::min((m + 255) / 256, h->sm_count * 8);
        row_stats_kernel<<<blocks, 256, 0, st>>>(m, rowptr, h->scratch_i);
        h->launches++;
        if (colidx && (h->opts[RSG_OPT_SPMV_ALGO] == RSG_SPMV_AUTO || h->opts[RSG_OPT_SPMV_ALGO] == RSG_SPMV_WINDOW || h->opts[RSG_OPT_SPMV_ALGO] == RSG_SPMV_TILE)) {
            const int groups = std::min((m + 31) / 32, 4096);
            scatter_stats_kernel<<<(groups * 32 + 255) / 256, 256, 0, st>>>(m, base, groups, rowptr, colidx, h->scratch_i);
            h->launches++;
        }
    } else {
        RSG_CUDA_TRY(cudaMemcpyAsync(h->scratch_i + 1, rowptr, sizeof(int), cudaMemcpyDeviceToDevice, st));
        RSG_CUDA_TRY(cudaMemcpyAsync(h->scratch_i + 2, rowptr, sizeof(int), cudaMemcpyDeviceToDevice, st));
    }
    RSG_CUDA_TRY(cudaMemcpyAsync(h->pinned_i, h->scratch_i, 8 * sizeof(int), cudaMemcpyDeviceToHost, st));
    RSG_CUDA_TRY(cudaStreamSynchronize(st));
    p.max_row = h->pinned_i[0];
    const double scatter = h->pinned_i[4] > 0 ? (double)h->pinned_i[3] / h->pinned_i[4] : 0.0; // x lines per row, neighbouring rows
    p.scattered = scatter > 0.5;
    p.first = h->pinned_i[1] - base;
    p.nnz = h->pinned_i[2] - h->pinned_i[1];
    if (p.nnz < 0 || p.first < 0) return RSG_BAD_ARG;

    int algo = h->opts[RSG_OPT_SPMV_ALGO];
    int variant = h->opts[RSG_OPT_SPMV_VARIANT];
    if (algo == RSG_SPMV_STREAM || algo == RSG_SPMV_PIPE) { algo = RSG_SPMV_TILE; variant = 0; } // retired kernels
    const bool aligned = (((uintptr_t)colidx) & 15) == 0;
    const bool auto_algo = algo == RSG_SPMV_AUTO;
    const double avg = (double)p.nnz / std::max(m, 1);
    bool try_window = algo == RSG_SPMV_WINDOW;
    if (algo == RSG_SPMV_WINDOW) { algo = RSG_SPMV_TILE; variant = avg < 12.0 ? 5 : (vbytes >= 8 ? 0 : 1); } // what a declined window plan becomes
    if (auto_algo) {
        // A row reduced by one CTA streams at one SM's share of the bandwidth (~1/148); past
        // roughly nnz/64 entries that single row outlasts the rest of the product: merge-path.
        const int cap = RowBlockShape<4096>::CAP;
        const long long skew = std::max<long long>(cap, (long long)p.nnz / 64);
        // aligned arrays: the TILE kernel, which cuts rows longer than a CTA's stage into one-CTA segments (up to
        // a few thousand such rows; beyond that pipe_build_plan declines and merge-path takes over)
        algo = aligned ? RSG_SPMV_TILE : (p.max_row > skew ? RSG_SPMV_MERGE : RSG_SPMV_ROWBLOCK);
        // measured (tools/spmv_sweep.py).  Config 2: ~24-28 KB of entries per CTA is the sweet spot: 8-byte values ->
        // 128 threads x 2304 entries, 4-byte values -> 128 threads x 4352 entries.  Short rows (5-point stencils,
        // power-law rows; configs 1 and 4) are bound by the latency of the x gathers, not by the stream: a small
        // stage (128 threads x 1152 entries, 15 KB) keeps ~60 warps resident per SM
        if (algo == RSG_SPMV_TILE) {
            variant = avg < 12.0 ? 5 : (vbytes >= 8 ? 0 : 1);
            // Neighbouring rows that read neighbouring x (stencil / FEM / banded orderings: scatter statistic ~0.1):
            // the WINDOW kernel -- the tile's x window travels to shared memory with the tile, no gather is left on the
            // critical path (spmv_window.cu).  It declines patterns without compact windows.
            if (scatter <= 0.5 && m >= 64) try_window = true;
            // Two cases go to the plain sub-warp-per-row kernel instead (batched loads, no shared memory, so all 228 KB
            // of the SM stay L1 for x; profiles/r1_vector_lanes_sweep.txt, profiles/r1_auto_vs_vector.txt):
            //  * rows whose neighbours gather from unrelated x lines (hashed / badly ordered columns): thread-per-row
            //    gathers out of a staged tile no longer coalesce and L1 capacity for x is what helps (1.7-2.5x on the
            //    banded stand-ins of the 36-matrix sweep);
            //  * matrices whose entries fit in half of L2 (resident after the first product of a solver loop) and that the
            //    window kernel declined: the tile kernel's per-CTA set-up is a visible part of a 10-20 us product.
            // Not with dense rows (a lane walks its share of the row serially: the kernel widens to at most 32 lanes)
            // and not when tiles must wait on halo flags.
            const int vl = vector_lanes(avg);
            const bool no_halo = h->opts[RSG_OPT_SPMV_HALO_START] == 0;
            const int longest_ok = (scatter > 0.5 && avg >= 16.0) ? 4096 : 128 * vl;
            const bool l2_resident = (long long)p.nnz * (vbytes + 4) < (long long)h->l2_bytes / 2;
            if (no_halo && (scatter > 0.5 || l2_resident) && p.max_row <= longest_ok) { algo = RSG_SPMV_VECTOR; variant = 0; }
        }
    }
    if (try_window && aligned && m > 0 && p.nnz > 0) {
        SpmvPlan wp = p;
        wp.algo = RSG_SPMV_WINDOW;
        wp.variant = h->opts[RSG_OPT_SPMV_ALGO] == RSG_SPMV_WINDOW ? h->opts[RSG_OPT_SPMV_VARIANT] : 0;
        const int rc = window_build_plan(h, wp, vbytes, xbytes, auto_algo);
        if (rc == RSG_OK) {
            RSG_CUDA_TRY(cudaGetLastError());
            h->spmv_plans.push_back(wp);
            *out = &h->spmv_plans.back();
            return RSG_OK;
        }
        cudaFree(wp.wdesc); cudaFree(wp.widx); cudaFree(wp.rp0);
        (void)cudaGetLastError();
        if (rc != RSG_UNSUPPORTED) return rc;
    }
    if (algo == RSG_SPMV_TILE && !aligned) algo = RSG_SPMV_ROWBLOCK; // bulk copies need 16-byte aligned arrays
    const int nvar = algo == RSG_SPMV_TILE ? tile_num_variants() : kNumVariants;
    if (variant < 0 || variant >= nvar) variant = 0;
    p.variant = variant;
    p.tile = kVariants[algo == RSG_SPMV_TILE ? 0 : variant].tile;
    p.algo = algo;
    const long long items = (long long)m + p.nnz;
    if (algo == RSG_SPMV_TILE && m > 0) {
        const SpmvPlan bare = p; // nothing allocated yet
        const int rc = pipe_build_plan(h, p, auto_algo);
        if (rc == RSG_UNSUPPORTED && auto_algo) { // too many long rows to split: merge-path
            spmv_free_plan(p);
            p = bare;
            algo = p.algo = RSG_SPMV_MERGE;
            p.variant = 0;
            p.tile = kVariants[0].tile;
        } else {
            RSG_TRY(rc);
        }
    }
    if (algo == RSG_SPMV_ROWBLOCK && m > 0) {
        p.nblk = (int)((items + p.tile - 1) / p.tile);
        RSG_CUDA_TRY(cudaMalloc(&p.blk_row, sizeof(int) * ((size_t)p.nblk + 1)));
        p.bytes += sizeof(int) * ((size_t)p.nblk + 1);
        rowblock_partition_kernel<<<(p.nblk + 1 + 255) / 256, 256, 0, st>>>(m, rowptr, p.tile, p.nblk, p.blk_row);
        h->launches++;
    } else if (algo == RSG_SPMV_MERGE && m > 0) {
        p.merge_items = kMergeThreads * kMergeIpt;
        p.ntiles = (int)((items + p.merge_items - 1) / p.merge_items);
        RSG_CUDA_TRY(cudaMalloc(&p.coords, sizeof(int2) * ((size_t)p.ntiles + 1)));
        RSG_CUDA_TRY(cudaMalloc(&p.carry_row, sizeof(int) * (size_t)p.ntiles));
        RSG_CUDA_TRY(cudaMalloc(&p.carry_val, sizeof(double) * (size_t)p.ntiles));
        p.bytes += (sizeof(int2) + sizeof(int) + sizeof(double)) * ((size_t)p.ntiles + 1);
        merge_coords_kernel<<<(p.ntiles + 1 + 255) / 256, 256, 0, st>>>(m, p.nnz, base, rowptr, p.merge_items, p.ntiles, p.coords);
        h->launches++;
    }
    RSG_CUDA_TRY(cudaGetLastError());
    h->spmv_plans.push_back(p);
    *out = &h->spmv_plans.back();
    return RSG_OK;
}

template <typename TV, typename TX, int THREADS, int TILE, int HINT, bool VEC>
static int launch_rowblock_one(rsg_context *h, const SpmvPlan *p, const TV *values, TX alpha, const TX *x, TX beta, TX *y)
{
    auto kern = spmv_rowblock_kernel<TV, TX, THREADS, TILE, HINT, VEC>;
    const size_t smem = sizeof(TX) * (RowBlockShape<TILE>::CAP + 4) + sizeof(int) * (TILE + 1);
    static PerDeviceOnce configured; // function attributes are per device (and per instantiation)
    if (configured.need(h->device)) {
        RSG_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    }
    kern<<<p->nblk, THREADS, smem, h->stream>>>(p->base, p->rowptr, p->colidx, values, x, y, alpha, beta, p->blk_row,
                                               h->opts[RSG_OPT_SPMV_REDUCE]);
    h->launches++;
    return RSG_OK;
}

template <typename TV, typename TX, int THREADS, int TILE>
static int launch_rowblock_shape(rsg_context *h, const SpmvPlan *p, bool vec, const TV *values, TX alpha, const TX *x, TX beta, TX *y)
{
    const int hint = h->opts[RSG_OPT_SPMV_CACHE_HINT];
    if (!vec) return launch_rowblock_one<TV, TX, THREADS, TILE, 0, false>(h, p, values, alpha, x, beta, y);
    switch (hint) {
    case 0: return launch_rowblock_one<TV, TX, THREADS, TILE, 0, true>(h, p, values, alpha, x, beta, y);
    case 2: return launch_rowblock_one<TV, TX, THREADS, TILE, 2, true>(h, p, values, alpha, x, beta, y);
    default: return launch_rowblock_one<TV, TX, THREADS, TILE, 1, true>(h, p, values, alpha, x, beta, y);
    }
}

template <typename TV, typename TX>
static int launch_rowblock(rsg_context *h, const SpmvPlan *p, bool vec, const TV *values, TX alpha, const TX *x, TX beta, TX *y)
{
    switch (p->variant) {
    case 1: return launch_rowblock_shape<TV, TX, 128, 2048>(h, p, vec, values, alpha, x, beta, y);
    case 2: return launch_rowblock_shape<TV, TX, 256, 2048>(h, p, vec, values, alpha, x, beta, y);
    case 3: return launch_rowblock_shape<TV, TX, 512, 8192>(h, p, vec, values, alpha, x, beta, y);
    case 4: return launch_rowblock_shape<TV, TX, 128, 1024>(h, p, vec, values, alpha, x, beta, y);
    default: return launch_rowblock_shape<TV, TX, 256, 4096>(h, p, vec, values, alpha, x, beta, y);
    }
}

template <typename TV, typename TX>
static int launch_merge(rsg_context *h, const SpmvPlan *p, bool vec, const TV *values, TX alpha, const TX *x, TX beta, TX *y)
{
    const int hint = h->opts[RSG_OPT_SPMV_CACHE_HINT];
    TX *cv = reinterpret_cast<TX *>(p->carry_val);
#define RSG_MERGE(H, V)                                                                                          \
    spmv_merge_kernel<TV, TX, 256, 8, H, V><<<p->ntiles, 256, 0, h->stream>>>(p->m, p->base, p->rowptr, p->colidx, \
                                                                              values, x, y, alpha, beta, p->coords, \
                                                                              p->carry_row, cv)
    if (!vec) RSG_MERGE(0, false);
    else if (hint == 0) RSG_MERGE(0, true);
    else if (hint == 2) RSG_MERGE(2, true);
    else RSG_MERGE(1, true);
#undef RSG_MERGE
    merge_fixup_kernel<TX><<<(p->ntiles + 255) / 256, 256, 0, h->stream>>>(p->ntiles, p->carry_row, cv, y, alpha);
    h->launches += 2;
    return RSG_OK;
}

static int set_x_persistence(rsg_context *h, const void *x, size_t bytes)
{
    if (!h->opts[RSG_OPT_SPMV_X_PERSIST]) {
        if (h->persist_ptr) {
            cudaStreamAttrValue attr = {};
            attr.accessPolicyWindow.num_bytes = 0;
            cudaStreamSetAttribute(h->stream, cudaStreamAttributeAccessPolicyWindow, &attr);
            h->persist_ptr = nullptr; h->persist_bytes = 0;
        }
        return RSG_OK;
    }
    if (h->persist_ptr == x && h->persist_bytes == bytes) return RSG_OK;
    cudaDeviceProp prop;
    RSG_CUDA_TRY(cudaGetDeviceProperties(&prop, h->device));
    const size_t win = std::min(bytes, (size_t)prop.accessPolicyMaxWindowSize);
    if (win == 0) return RSG_OK;
    cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, std::min((size_t)prop.persistingL2CacheMaxSize, win));
    cudaStreamAttrValue attr = {};
    attr.accessPolicyWindow.base_ptr = const_cast<void *>(x);
    attr.accessPolicyWindow.num_bytes = win;
    attr.accessPolicyWindow.hitRatio = 1.0f;
    attr.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
    attr.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
    RSG_CUDA_TRY(cudaStreamSetAttribute(h->stream, cudaStreamAttributeAccessPolicyWindow, &attr));
    h->persist_ptr = x; h->persist_bytes = bytes;
    return RSG_OK;
}

template <typename TV, typename TX>
int spmv_run(rsg_context *h, int m, int n, int nnz_hint, int base, const int *rowptr, const int *colidx,
             const TV *values, TX alpha, const TX *x, TX beta, TX *y)
{
    if (m < 0 || n < 0 || (base != 0 && base != 1) || !rowptr || (m > 0 && !y)) return RSG_BAD_ARG;
    if (m == 0) return RSG_OK;
    SpmvPlan *p = nullptr;
    constexpr int pkey = (int)sizeof(TV) | ((int)sizeof(TX) << 4);
    RSG_TRY(spmv_get_plan(h, m, n, nnz_hint, base, rowptr, colidx, false, &p, pkey));
    if (p->nnz > 0 && (!colidx || !values || !x)) return RSG_BAD_ARG;
    if (p->algo == RSG_SPMV_WINDOW && ((((uintptr_t)x) | ((uintptr_t)values)) & 15) != 0) {
        // the x segments and the values travel by 16-byte bulk copies: an odd view of x takes the tile / row-block plan
        const int keep_algo = h->opts[RSG_OPT_SPMV_ALGO], keep_var = h->opts[RSG_OPT_SPMV_VARIANT];
        h->opts[RSG_OPT_SPMV_ALGO] = ((((uintptr_t)values) & 15) == 0) ? RSG_SPMV_TILE : RSG_SPMV_ROWBLOCK;
        h->opts[RSG_OPT_SPMV_VARIANT] = h->opts[RSG_OPT_SPMV_ALGO] == RSG_SPMV_TILE ? (p->nnz < 12LL * m ? 5 : (sizeof(TV) >= 8 ? 0 : 1)) : 0;
        const int rc = spmv_get_plan(h, m, n, nnz_hint, base, rowptr, colidx, false, &p, pkey);
        h->opts[RSG_OPT_SPMV_ALGO] = keep_algo; h->opts[RSG_OPT_SPMV_VARIANT] = keep_var;
        RSG_TRY(rc);
    }
    h->last_spmv_algo = p->algo;
    h->last_spmv_maxrow = p->max_row;
    RSG_TRY(set_x_persistence(h, x, sizeof(TX) * (size_t)n));
    const bool vec = (((uintptr_t)colidx | (uintptr_t)values) & 15) == 0;
    if ((h->wait_flags || h->fuse.ctas) && p->algo != RSG_SPMV_TILE && p->algo != RSG_SPMV_WINDOW) { h->wait_flags = nullptr; h->fuse.ctas = 0; h->chain.npeers = 0; return RSG_UNSUPPORTED; }
    if (h->packets.pkts && p->algo != RSG_SPMV_WINDOW) { h->packets.pkts = nullptr; h->wait_flags = nullptr; h->fuse.ctas = 0; return RSG_UNSUPPORTED; }
    if (h->chain.npeers > 0 && (p->algo != RSG_SPMV_WINDOW || p->wprod <= 0)) { h->wait_flags = nullptr; h->chain.npeers = 0; return RSG_UNSUPPORTED; }
    switch (p->algo) {
    case RSG_SPMV_ROWBLOCK:
        h->last_spmv_blocks = p->nblk;
        RSG_TRY((launch_rowblock<TV, TX>(h, p, vec, values, alpha, x, beta, y)));
        break;
    case RSG_SPMV_TILE:
        if ((((uintptr_t)values) & 15) != 0) return RSG_BAD_ARG; // plan was made for aligned arrays
        RSG_TRY((tile_launch<TV, TX>(h, p, values, alpha, x, beta, y)));
        break;
    case RSG_SPMV_WINDOW:
        RSG_TRY((window_launch<TV, TX>(h, p, values, alpha, x, beta, y)));
        break;
    case RSG_SPMV_MERGE:
        h->last_spmv_blocks = p->ntiles;
        RSG_TRY((launch_merge<TV, TX>(h, p, vec, values, alpha, x, beta, y)));
        break;
    case RSG_SPMV_VECTOR: {
        const double avg = (double)p->nnz / std::max(m, 1);
        const int threads = 256;
#define RSG_VEC_B(G, B)                                                                                        \
    spmv_vector_kernel<TV, TX, G, B><<<(int)(((long long)m * G + threads - 1) / threads), threads, 0, h->stream>>>( \
        m, base, rowptr, colidx, values, x, y, alpha, beta)
#define RSG_VEC(G)                                                                                             \
    do {                                                                                                       \
        if (vb == 8) RSG_VEC_B(G, 8); else if (vb == 2) RSG_VEC_B(G, 2); else RSG_VEC_B(G, 4);                  \
    } while (0)
        // entries per lane and batch: 8 with 4-byte x (45.0 vs 51.6 us on config 2), 4 with 8-byte x (64.6 vs 67.8 us);
        // RSG_OPT_SPMV_VARIANT 1 / 2 / 3 force 8 / 2 / 4 (profiles/r1_vector_batch_lanes.txt)
        int lanes = h->opts[RSG_OPT_SPMV_REDUCE]; // 0 = by row lengths; a power of two forces it (A/B)
        if (lanes != 1 && lanes != 2 && lanes != 4 && lanes != 8 && lanes != 16 && lanes != 32) {
            lanes = vector_lanes(avg);
            while (lanes < 32 && p->max_row > 128 * lanes) lanes *= 2; // the longest row bounds a lane's serial walk
        }
        // a batch of 8 only pays when a lane has more than 4 entries to take (5-point rows, fp32: 16.4 us with 4, 18.5 with 8)
        const int vb = p->variant == 1 ? 8 : p->variant == 2 ? 2 : p->variant == 3 ? 4 : (sizeof(TX) == 4 && avg > 4.0 * lanes ? 8 : 4);
        switch (lanes) {
        case 1: RSG_VEC(1); break;
        case 2: RSG_VEC(2); break;
        case 4: RSG_VEC(4); break;
        case 8: RSG_VEC(8); break;
        case 16: RSG_VEC(16); break;
        default: RSG_VEC(32); break;
        }
#undef RSG_VEC_B
#undef RSG_VEC
        h->last_spmv_blocks = 0;
        h->launches++;
        break;
    }
    case RSG_SPMV_SCALAR:
        spmv_scalar_kernel<TV, TX><<<(m + 255) / 256, 256, 0, h->stream>>>(m, base, rowptr, colidx, values, x, y, alpha, beta);
        h->last_spmv_blocks = (m + 255) / 256;
        h->launches++;
        break;
    default:
        return RSG_BAD_ARG;
    }
    RSG_CUDA_TRY(cudaGetLastError());
    return RSG_OK;
}

template int spmv_run<double, double>(rsg_context *, int, int, int, int, const int *, const int *, const double *, double,
                                      const double *, double, double *);
template int spmv_run<float, float>(rsg_context *, int, int, int, int, const int *, const int *, const float *, float,
                                    const float *, float, float *);
template int spmv_run<float, double>(rsg_context *, int, int, int, int, const int *, const int *, const float *, double,
                                     const double *, double, double *);

int row_partition(rsg_context *h, int m, int base, const int *rowptr, int ngpu, int balanced, int *row_starts_host)
{
    if (ngpu < 1 || ngpu > 63 || m < 0 || !rowptr || !row_starts_host) return RSG_BAD_ARG;
    row_partition_kernel<<<1, 64, 0, h->stream>>>(m, base, rowptr, ngpu, balanced, h->scratch_i);
    h->launches++;
    RSG_CUDA_TRY(cudaMemcpyAsync(h->pinned_i, h->scratch_i, sizeof(int) * (size_t)(ngpu + 1), cudaMemcpyDeviceToHost, h->stream));
    RSG_CUDA_TRY(cudaStreamSynchronize(h->stream));
    for (int g = 0; g <= ngpu; ++g) row_starts_host[g] = h->pinned_i[g];
    return RSG_OK;
}

} // namespace rsg
