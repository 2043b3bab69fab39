// spmv_tile.cu -- the general TMA-staged SpMV kernel (RSG_SPMV_TILE): one tile per CTA, many small CTAs per SM.
// Replaces the cusparseSpMV call of the reference harness (/root/reference/GPU/spmv.cu:174-186) for every
// matrix the x-window kernel (spmv_window.cu) declines: scattered columns, rows longer than a stage.
//
// One elected thread issues three 1-D bulk async copies (cp.async.bulk -> UBLKCP): the row-pointer slice,
// colidx and values, all completing on one mbarrier; the CTA then reduces its rows (thread-, sub-warp- or
// warp-per-row, x gathered in row order so that neighbouring rows of banded matrices read neighbouring x).
//
// Plan (built on device by pipe_build_plan): a zero-based, padded copy of the row pointers (so that any slice
// can be bulk-copied from a 16-byte aligned address) and one descriptor per tile (first row, first entry,
// longest row).  Tiles that do not fit a stage take a synchronous multi-pass path; a row longer than a whole
// stage is cut into segments with a CTA each (LongSplit).
// (Round 1 also carried a persistent multi-stage variant of this kernel, RSG_SPMV_PIPE, and a first TMA
// version, RSG_SPMV_STREAM; both were measured slower on every config -- DESIGN.md 5.3 -- and are retired:
// the enum values now select this kernel.)
#include "spmv_device.cuh"

#include <cstring>

namespace rsg {

template <typename TV, int CAPE, int ROWS> struct PipeStage {
    static constexpr size_t val_off = 0;
    static constexpr size_t col_off = sizeof(TV) * (size_t)(CAPE + 8);
    static constexpr size_t rp_off = col_off + sizeof(int) * (size_t)(CAPE + 8);
    static constexpr size_t bytes = (rp_off + sizeof(int) * (size_t)(ROWS + 8) + 127) & ~(size_t)127;
};

struct PipeMeta { // written by the producer, read by everybody after the mbarrier wait
    int r0, nr, k0, k1; // rows [r0,r0+nr), zero-based entries [k0,k1)
    int maxlen, staged_end; // entries [k0&~3, staged_end) arrived by bulk copy (0: nothing staged)
    int rp_shift, pad;      // s_rp[rp_shift + i] = zero-based start of row r0+i
};

// The rows of one tile out of its stage: any row lengths, multi-pass when the tile does not fit one stage,
// CTA-per-row for a row longer than a stage (unless the segment CTAs take it).  COH: coherent x gathers.
template <typename TV, typename TX, int NC, int CAPE, typename Stage, bool COH>
__device__ __forceinline__ void tile_reduce(unsigned char *sp, const PipeMeta &mt, int base, const int *__restrict__ colidx,
                                            const TV *__restrict__ values, const TX *__restrict__ x, TX *__restrict__ y,
                                            TX alpha, TX beta, int reduce_opt, uint64_t spol, uint64_t *s_aux, bool skip_long)
{
    const int tid = threadIdx.x;
    TV *s_val = reinterpret_cast<TV *>(sp + Stage::val_off);
    int *s_col = reinterpret_cast<int *>(sp + Stage::col_off);
    const int *s_rp = reinterpret_cast<const int *>(sp + Stage::rp_off) + mt.rp_shift;
    const int nr = mt.nr, r0 = mt.r0;
    uint32_t aux_parity = 0;
    int cur = 0;
    bool staged = mt.staged_end > 0;
    while (cur < nr) {
        const int k0 = s_rp[cur];
        const int sb = k0 & ~3;
        int pe, k1;
        if (staged) {
            pe = nr; k1 = mt.k1;
            if (mt.staged_end < k1) { // only the last tile of a matrix whose entry count is not a multiple of 4
                const int k = mt.staged_end + tid;
                if (tid < 4 && k < k1) { s_val[k - sb] = values[k]; s_col[k - sb] = colidx[k]; }
                __syncthreads();
            }
        } else { // synchronous multi-pass path for a tile that does not fit one stage
            if (s_rp[nr] - sb <= CAPE) {
                pe = nr;
            } else { // largest pe in [cur,nr] with s_rp[pe]-sb <= CAPE
                int lo = cur, hi = nr;
                while (lo < hi) {
                    const int mid = (lo + hi + 1) >> 1;
                    if (s_rp[mid] - sb <= CAPE) lo = mid; else hi = mid - 1;
                }
                pe = lo;
            }
            if (pe == cur) { // this row alone exceeds a stage: CTA-per-row from global memory
                if (skip_long && s_rp[cur + 1] - k0 > CAPE) { cur += 1; continue; } // reduced by the segment CTAs (LongSplit)
                const TX tot = cta_row_sum<TV, TX, NC, 1, true, 0, COH>(k0, s_rp[cur + 1], colidx, values, x, base,
                                                                        reinterpret_cast<TX *>(s_val), spol);
                if (tid == 0) store_y(y, r0 + cur, tot, alpha, beta);
                cur += 1;
                continue;
            }
            k1 = s_rp[pe];
            const int qe = k1 & ~3;
            if (tid == 0 && qe > sb) {
                const uint32_t nq = (uint32_t)(qe - sb);
                mbar_expect_tx(s_aux, nq * (uint32_t)(sizeof(TV) + sizeof(int)));
                bulk_g2s(s_val, values + sb, nq * (uint32_t)sizeof(TV), s_aux, spol);
                bulk_g2s(s_col, colidx + sb, nq * 4u, s_aux, spol);
            }
            {
                const int k = max(qe, sb) + tid;
                if (tid < 4 && k < k1) { s_val[k - sb] = values[k]; s_col[k - sb] = colidx[k]; }
            }
            __syncthreads();
            if (qe > sb) { mbar_wait(s_aux, aux_parity); aux_parity ^= 1; }
        }
        const int nrows = pe - cur;
        int g = reduce_opt & 0xff;
        const bool prefer2 = (reduce_opt >> 16) & 1; // plan: about threads/2 rows per tile
        if (g == 0) { // lanes per row from the tile's row-length statistics
            const int avg = (k1 - k0) / nrows;
            const int len = max(avg, mt.maxlen >> 3);
            g = len <= 40 ? 1 : len <= 80 ? 4 : len <= 160 ? 8 : len <= 320 ? 16 : 32;
            while (g < 32 && nrows * g * 2 <= NC) g <<= 1; // few rows: spread them over the CTA
            if (g == 2) g = (avg <= 40 && mt.maxlen <= 64) ? 1 : 4;
            if (prefer2 && g == 1 && mt.maxlen <= 64 && nrows * 2 <= NC) g = 2;
        }
        const int *rp = s_rp + cur;
        switch (g) {
        case 1: { // thread per row: unpredicated whole batches (8 gathers per batch for fp64 x, 16 for fp32)
            constexpr int BB = sizeof(TX) == 4 ? 16 : 8;
            if (sizeof(TV) == sizeof(TX) && (reduce_opt & 0xff) == 0 && ((reduce_opt >> 17) & 1)) {
                // scattered columns (plan: SpmvPlan::scatter): one round of gathers for the whole tile, products staged in place
                if constexpr (sizeof(TV) == sizeof(TX)) {
                    stage_products<TV, TX, NC, COH>(s_val, s_col, rp[0] - sb, rp[nrows] - sb, x, base);
                    __syncthreads();
                    sum_staged_rows<TV, TX, NC, 8, 32>(s_val, rp, sb, nrows, r0 + cur, y, alpha, beta, mt.maxlen > 32);
                }
                break;
            }
            if ((reduce_opt & 0xff) == 0 && mt.maxlen > 2 * BB) {
                // chosen by the tile's statistics (not forced): rows of more than two batches go to eight lanes each afterwards
                reduce_rows_lean2<TV, TX, NC, BB, COH, 2 * BB>(s_val, s_col, rp, sb, nrows, r0 + cur, x, base, y, alpha, beta);
                reduce_rows_long<TV, TX, NC, 8, BB, COH, 2 * BB>(s_val, s_col, rp, sb, nrows, r0 + cur, x, base, y, alpha, beta);
            } else {
                reduce_rows_lean2<TV, TX, NC, BB, COH>(s_val, s_col, rp, sb, nrows, r0 + cur, x, base, y, alpha, beta);
            }
            break;
        }
        case 2: reduce_rows_lean2g<TV, TX, NC, 2, 8, COH>(s_val, s_col, rp, sb, nrows, r0 + cur, x, base, y, alpha, beta); break;
        case 4: reduce_rows_lean<TV, TX, NC, 4, COH>(s_val, s_col, rp, sb, nrows, r0 + cur, x, base, y, alpha, beta); break;
        case 8: reduce_rows_lean<TV, TX, NC, 8, COH>(s_val, s_col, rp, sb, nrows, r0 + cur, x, base, y, alpha, beta); break;
        case 16: reduce_rows_lean<TV, TX, NC, 16, COH>(s_val, s_col, rp, sb, nrows, r0 + cur, x, base, y, alpha, beta); break;
        default: reduce_rows_lean<TV, TX, NC, 32, COH>(s_val, s_col, rp, sb, nrows, r0 + cur, x, base, y, alpha, beta); break;
        }
        cur = pe;
        if (!staged || cur < nr) __syncthreads(); // the synchronous path reuses the buffers
        staged = false;
    }
}

// One tile per CTA: latency is hidden by many small co-resident CTAs.  MINB states the residency the shared-memory
// footprint allows anyway (7 CTAs of 128 threads -> up to 72 registers), so ptxas no longer squeezes the kernel
// into 40 registers with a spill (round 1: 16-byte stack frame, STL in the prologue).
// Critical path of a CTA: descriptor load -> three bulk copies (row pointers, colidx, values) -> mbarrier wait ->
// reduction.  COH = 1 is the multi-GPU instantiation: tiles marked as halo readers (and the long-row segment CTAs)
// await the peers' arrival flags first and gather x with coherent loads.
template <typename TV, typename TX, int THREADS, int CAPE, int ROWS, int MINB, bool COH>
__global__ void __launch_bounds__(THREADS, MINB)
spmv_tile_kernel(int base, const int *__restrict__ rp0, const int *__restrict__ colidx, const TV *__restrict__ values,
                 const TX *__restrict__ x, TX *__restrict__ y, TX alpha, TX beta, const int4 *__restrict__ desc,
                 int nnz_end, int ntiles, int reduce_opt, int early_values, const int *wait_flags, int wait_n, int wait_epoch,
                 int *err_flag, int max_spins, const HaloFuse fuse, const LongSplit ls)
{
    typedef PipeStage<TV, CAPE, ROWS> Stage;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    __shared__ __align__(8) uint64_t s_full;
    __shared__ __align__(8) uint64_t s_aux;
    const int tid = threadIdx.x;
    // programmatic dependent launch, as in the window kernel: the next kernel of the stream may start once every CTA of
    // this grid is here; this grid reads only pattern data (descriptor, row pointers, colidx -- the plan is keyed on the
    // pattern) and, if promised static, the values before grid_dependency_wait()
    if (!(COH && fuse.pull) || (int)blockIdx.x < fuse.ctas) grid_launch_dependents(); // (pull exchange: see spmv_window.cu)
    unsigned long long *tr = ls.trace ? ls.trace + 4 * (size_t)blockIdx.x : nullptr;
    if (tr && tid == 0) tr[0] = global_ns();
    if (COH && (int)blockIdx.x < fuse.ctas) {
        // ---- fused exchange: these CTAs are dispatched first; they publish this rank's boundary entries of x
        // to the peers (plain stores into P2P-mapped memory over NVLink) while the tiles behind them multiply
        if (fuse.pull) { halo_pull<TX, THREADS>(fuse, err_flag, max_spins); return; }
        grid_dependency_wait();
        halo_fuse_push<TX, THREADS>(fuse, x);
        return;
    }
    const int seg_id = (int)blockIdx.x - (COH ? fuse.ctas : 0);
    if (seg_id < ls.nseg) {
        // ---- one segment of a long row, straight from global memory with 128-bit loads; the last segment of the
        // row to arrive adds the partial sums in segment order and stores y
        grid_dependency_wait();
        if (COH && fuse.pull) grid_launch_dependents();
        if (COH && wait_n > 0) { // a dense row may reach into the halo: its segments wait like a marked tile
            halo_flag_wait(wait_flags, wait_n, wait_epoch, err_flag, max_spins);
            __syncthreads();
        }
        if (tr && tid == 0) tr[1] = global_ns();
        const int4 sg = ls.seg[seg_id];
        const TX tot = cta_row_sum<TV, TX, THREADS, 1, true, 0, COH>(sg.y, sg.z, colidx, values, x, base, reinterpret_cast<TX *>(smem_raw),
                                                                     make_evict_first_policy());
        if (tr && tid == 0) tr[2] = global_ns();
        __shared__ int s_lastseg;
        if (tid == 0) {
            const int2 lr = ls.lrow[sg.w];
            *reinterpret_cast<volatile TX *>(ls.part + seg_id) = tot; // one 8-byte slot per segment, whatever TX is
            __threadfence();
            s_lastseg = atomicAdd(ls.cnt + sg.w, 1) == lr.y - 1;
        }
        __syncthreads();
        if (s_lastseg) {
            // the last segment of the row to arrive adds the partial sums -- the whole CTA, in an order fixed by the segment
            // count alone (thread t takes partials t, t + THREADS, ...; then lanes, then warps): deterministic, and one round of
            // loads instead of a thread walking ~100 partials one L2 round trip at a time (measured: up to 5-8 us, the end of
            // the product on config 4)
            __threadfence();
            const int2 lr = ls.lrow[sg.w];
            TX acc = 0;
            for (int i = tid; i < lr.y; i += THREADS) acc += *reinterpret_cast<volatile TX *>(ls.part + lr.x + i);
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) acc += __shfl_down_sync(0xffffffffu, acc, o);
            TX *red = reinterpret_cast<TX *>(smem_raw);
            if ((tid & 31) == 0) red[tid >> 5] = acc;
            __syncthreads();
            if (tid == 0) {
                TX sum = 0;
                for (int wv = 0; wv < THREADS / 32; ++wv) sum += red[wv];
                store_y(y, sg.x, sum, alpha, beta);
                ls.cnt[sg.w] = 0;
            }
        }
        if (tr && tid == 0) tr[3] = global_ns() | (1ull << 63);
        return;
    }
    const int tile_id = seg_id - ls.nseg;
    const int4 d0 = desc[tile_id], d1 = desc[tile_id + 1];
    if (tid == 0) { mbar_init(&s_full, 1); mbar_init(&s_aux, 1); }
    __syncthreads();
    const uint64_t spol = make_evict_first_policy();
    PipeMeta mt;
    mt.r0 = d0.x; mt.nr = d1.x - d0.x; mt.k0 = d0.y; mt.k1 = d1.y; mt.maxlen = d0.z; mt.pad = 0;
    // the last tile of a product that was armed with flags always awaits them, marked or not: a rank that feeds its
    // neighbours but reads nothing from them can then never run more than one step ahead of them
    const bool must_wait = COH && wait_n > 0 && (d0.w != 0 || tile_id == ntiles - 1);
    if (mt.nr <= 0 && !must_wait) return;
    const int ra = mt.r0 & ~3;
    mt.rp_shift = mt.r0 - ra;
    const uint32_t nrp = (uint32_t)((mt.rp_shift + mt.nr + 1 + 3) & ~3);
    const int sb = mt.k0 & ~3;
    int qe = min((mt.k1 + 3) & ~3, nnz_end & ~3);
    const bool fits = (mt.k1 - sb) <= CAPE && qe > sb;
    if (!fits) qe = sb;
    mt.staged_end = fits ? qe : 0;
    if (tid == 0 && mt.nr > 0) {
        const uint32_t nq = (uint32_t)(qe - sb);
        mbar_expect_tx(&s_full, nrp * 4u + nq * (uint32_t)(sizeof(TV) + sizeof(int)));
        bulk_g2s(smem_raw + Stage::rp_off, rp0 + ra, nrp * 4u, &s_full, spol);
        if (nq) {
            bulk_g2s(smem_raw + Stage::col_off, colidx + sb, nq * 4u, &s_full, spol);
            if (early_values) bulk_g2s(smem_raw + Stage::val_off, values + sb, nq * (uint32_t)sizeof(TV), &s_full, spol);
        }
    }
    grid_dependency_wait(); // x, y and (unless promised static) the values: what earlier work of the stream may have written
    if (COH && fuse.pull) grid_launch_dependents();
    if (tr && tid == 0) tr[1] = global_ns();
    if (tid == 0 && mt.nr > 0 && qe > sb && !early_values)
        bulk_g2s(smem_raw + Stage::val_off, values + sb, (uint32_t)(qe - sb) * (uint32_t)sizeof(TV), &s_full, spol);
    if (must_wait) {
        // this tile gathers x entries that peers write over NVLink: wait for their arrival flags
        // (while the bulk copies of the tile are in flight); every other tile never looks at them
        halo_flag_wait(wait_flags, wait_n, wait_epoch, err_flag, max_spins);
        __syncthreads();
        if (mt.nr <= 0) return;
    }
    mbar_wait(&s_full, 0);
    if (tr && tid == 0) tr[2] = global_ns();
    tile_reduce<TV, TX, THREADS, CAPE, Stage, COH>(smem_raw, mt, base, colidx, values, x, y, alpha, beta, reduce_opt, spol, &s_aux,
                                                   ls.nseg > 0);
    if (tr && tid == 0) tr[3] = global_ns();
}

// rows longer than `cape` -> list[] = (row, zero-based start, zero-based end, 0), count in *counter (may exceed cap)
__global__ void long_rows_kernel(int m, int base, int cape, const int *__restrict__ rowptr, int4 *list, int cap, int *counter)
{
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= m) return;
    const int a = rowptr[r], b = rowptr[r + 1];
    if (b - a > cape) {
        const int slot = atomicAdd(counter, 1);
        if (slot < cap) list[slot] = make_int4(r, a - base, b - base, 0);
    }
}

// desc[b] = (first row, its zero-based entry offset, longest row of the tile, 0); item coordinate
// = entries + w*rows, so a tile holds at most tile/w rows and tile + (its last row) entries.
__global__ void pipe_partition_kernel(int m, int base, const int *__restrict__ rowptr, int tile, int w, int ntiles, int4 *desc)
{
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b > ntiles) return;
    int r = m;
    if (b < ntiles) {
        const int rp0 = rowptr[0];
        const long long target = (long long)b * tile;
        int lo = 0, hi = m;
        while (lo < hi) {
            const int mid = lo + ((hi - lo) >> 1);
            if ((long long)(rowptr[mid] - rp0) + (long long)w * mid < target) lo = mid + 1; else hi = mid;
        }
        r = lo;
    }
    desc[b] = make_int4(r, rowptr[r] - base, 0, 0);
}

// one warp per tile: longest row -> desc[b].z (rows longer than `skip` > 0 are reduced elsewhere and do not count)
__global__ void pipe_maxlen_kernel(int ntiles, int base, int skip, const int *__restrict__ rowptr, int4 *desc)
{
    const int b = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (b >= ntiles) return;
    const int r0 = desc[b].x, r1 = desc[b + 1].x;
    int ml = 0;
    for (int r = r0 + lane; r < r1; r += 32) {
        const int len = rowptr[r + 1] - rowptr[r];
        if (skip <= 0 || len <= skip) ml = max(ml, len);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) ml = max(ml, __shfl_xor_sync(0xffffffffu, ml, o));
    if (lane == 0) desc[b].z = ml;
}

// one warp per tile: does the tile read a column >= halo_start (an x entry filled by a peer)?  -> desc[b].w
__global__ void pipe_halo_mark_kernel(int ntiles, int base, int halo_start, const int *__restrict__ colidx, int4 *desc)
{
    const int b = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (b >= ntiles) return;
    const int k0 = desc[b].y, k1 = desc[b + 1].y;
    int hit = 0;
    for (int k = k0 + lane; k < k1; k += 32) hit |= (colidx[k] - base >= halo_start);
    hit = __any_sync(0xffffffffu, hit);
    if (lane == 0) desc[b].w = hit ? 1 : 0;
}

// zero-based copy of the row pointers, padded to a multiple of 4 (+4) with the end offset
__global__ void pipe_rowptr_copy_kernel(int m, int padded, int base, const int *__restrict__ rowptr, int *__restrict__ rp0)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= padded) return;
    rp0[i] = rowptr[min(i, m)] - base;
}

// ------------------------------------------------------------------ host side
struct TileVariant { int threads, cape, rows; };
static const TileVariant kTileVariants[] = { // threads, entries per CTA, rows per CTA
    {128, 2304, 512},  // 0: 8-byte values (config 2 fp64: ~24-28 KB of entries per CTA)
    {128, 4352, 512},  // 1: 4-byte values
    {64, 1152, 256},   // 2
    {256, 4352, 512},  // 3
    {256, 2304, 512},  // 4
    {128, 1152, 256},  // 5: short rows (gather-latency bound): small stages, 56-60 resident warps per SM
    {128, 2304, 512},  // 6: shape of 0 WITHOUT the residency hint (round 1's build; A/B of the register cap)
    {128, 4352, 512},  // 7: shape of 1 without it
    {128, 768, 128},   // 8: short rows, one row per thread: 16 CTAs of 10 KB per SM, one round of gathers per CTA
    {64, 512, 64},     // 9: the same with 64-thread CTAs
};
static const int kNumTileVariants = (int)(sizeof(kTileVariants) / sizeof(kTileVariants[0]));
int tile_num_variants() { return kNumTileVariants; }

static constexpr int kMaxLongRows = 4096;  // rows split into segments per plan
static constexpr int kSegEntries = 1024;   // entries per segment CTA (128 threads x 2 quads: one pass, both quads' loads in flight together)
static constexpr int kMaxSegments = 1 << 20;

// LongSplit of a TILE plan.  strict: RSG_UNSUPPORTED when there are too many long rows (the AUTO caller then
// takes merge-path); otherwise such a plan simply keeps its long rows CTA-per-row inside the tiles.
static int tile_build_split(rsg_context *h, SpmvPlan &p, int cape, bool strict)
{
    cudaStream_t st = h->stream;
    int4 *list = nullptr;
    RSG_CUDA_TRY(cudaMalloc(&list, sizeof(int4) * kMaxLongRows));
    int *counter = h->scratch_i + 8;
    RSG_CUDA_TRY(cudaMemsetAsync(counter, 0, sizeof(int), st));
    long_rows_kernel<<<(p.m + 255) / 256, 256, 0, st>>>(p.m, p.base, cape, p.rowptr, list, kMaxLongRows, counter);
    h->launches++;
    int nlong = 0;
    cudaError_t e = cudaMemcpyAsync(&nlong, counter, sizeof(int), cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    std::vector<int4> rows;
    if (e == cudaSuccess && nlong > 0 && nlong <= kMaxLongRows) {
        rows.resize((size_t)nlong);
        e = cudaMemcpy(rows.data(), list, sizeof(int4) * (size_t)nlong, cudaMemcpyDeviceToHost);
    }
    cudaFree(list);
    if (e != cudaSuccess) { (void)cudaGetLastError(); return RSG_CUDA_ERROR; }
    if (nlong > kMaxLongRows) return strict ? RSG_UNSUPPORTED : RSG_OK;
    if (nlong == 0) return RSG_OK;
    std::sort(rows.begin(), rows.end(), [](const int4 &a, const int4 &b) { return a.x < b.x; });
    std::vector<int4> seg;
    std::vector<int2> lrow((size_t)nlong);
    for (int i = 0; i < nlong; ++i) {
        const int k0 = rows[i].y, k1 = rows[i].z;
        lrow[i].x = (int)seg.size();
        int a = k0;
        while (a < k1) { // cuts on multiples of 4 entries, so that the 128-bit loads of every segment are aligned
            int b = (a + kSegEntries) & ~3;
            if (b >= k1 - 64) b = k1; // no crumbs
            seg.push_back(make_int4(rows[i].x, a, b, i));
            a = b;
        }
        lrow[i].y = (int)seg.size() - lrow[i].x;
        if (seg.size() > (size_t)kMaxSegments) return strict ? RSG_UNSUPPORTED : RSG_OK;
    }
    p.nseg = (int)seg.size();
    p.nlong = nlong;
    RSG_CUDA_TRY(cudaMalloc(&p.seg, sizeof(int4) * seg.size()));
    RSG_CUDA_TRY(cudaMalloc(&p.lrow, sizeof(int2) * lrow.size()));
    RSG_CUDA_TRY(cudaMalloc(&p.seg_part, sizeof(double) * seg.size()));
    RSG_CUDA_TRY(cudaMalloc(&p.seg_cnt, sizeof(int) * lrow.size()));
    RSG_CUDA_TRY(cudaMemcpy(p.seg, seg.data(), sizeof(int4) * seg.size(), cudaMemcpyHostToDevice));
    RSG_CUDA_TRY(cudaMemcpy(p.lrow, lrow.data(), sizeof(int2) * lrow.size(), cudaMemcpyHostToDevice));
    RSG_CUDA_TRY(cudaMemset(p.seg_cnt, 0, sizeof(int) * lrow.size()));
    p.bytes += (sizeof(int4) + sizeof(double)) * seg.size() + (sizeof(int2) + sizeof(int)) * lrow.size();
    return RSG_OK;
}

int pipe_build_plan(rsg_context *h, SpmvPlan &p, bool strict_split)
{
    cudaStream_t st = h->stream;
    const TileVariant &v = kTileVariants[p.variant];
    const int m = p.m;
    // tile: about one row per thread (rows_per_thread option scales it), within the stage capacity
    const int tmax = v.cape * 8 / 9;
    const int w = (tmax + v.rows - 1) / v.rows;
    const double avg = m > 0 ? (double)p.nnz / m : 0.0;
    int rpt = h->opts[RSG_OPT_SPMV_ROWS_PER_THREAD];
    if (rpt <= 0) rpt = 1;
    int lanes = h->opts[RSG_OPT_SPMV_REDUCE]; // forced lanes per row: size the tile for threads/lanes rows
    if (lanes < 1 || lanes > 32) lanes = 1;
    if (p.algo == RSG_SPMV_TILE && h->opts[RSG_OPT_SPMV_REDUCE] == 0 && avg >= 12.0 && h->opts[RSG_OPT_SPMV_X_HINT] != 4) { // X_HINT 4: A/B switch
        // rows the stage can hold vs threads of the CTA: config 2 in fp64 fits 66 rows of 27 entries under 128
        // threads -- with one thread per row half the CTA idles.  Two lanes per row (tile sized for threads/2 rows)
        // keeps every thread busy: 63.5 -> 61.4 us (tools/ab_option.py SPMV_REDUCE 0 2)
        const double rows_fit = tmax / (avg + w);
        if (v.threads / rows_fit >= 1.5 && v.threads / rows_fit < 3.0) { lanes = 2; p.lanes_hint = 2; }
    }
    long long want = (long long)((avg + w) * (v.threads / lanes) * rpt + 0.5);
    int tile = (int)std::min<long long>(tmax, std::max<long long>(want, 512));
    if (h->opts[RSG_OPT_SPMV_LONG_ROW] > 0) tile = std::min(tmax, std::max(64, h->opts[RSG_OPT_SPMV_LONG_ROW]));
    p.tile = tile;
    const long long items = (long long)w * m + p.nnz;
    p.nblk = (int)std::max<long long>(1, (items + tile - 1) / tile);
    p.rp0_len = ((m + 1 + 3) & ~3) + 4;
    RSG_CUDA_TRY(cudaMalloc(&p.pdesc, sizeof(int4) * ((size_t)p.nblk + 1)));
    RSG_CUDA_TRY(cudaMalloc(&p.rp0, sizeof(int) * (size_t)p.rp0_len));
    p.bytes += sizeof(int4) * ((size_t)p.nblk + 1) + sizeof(int) * (size_t)p.rp0_len;
    pipe_partition_kernel<<<(p.nblk + 1 + 255) / 256, 256, 0, st>>>(m, p.base, p.rowptr, tile, w, p.nblk, p.pdesc);
    // long rows first: the lanes-per-row choice of a tile must not see a row its CTA will skip
    if (p.algo == RSG_SPMV_TILE && p.max_row > v.cape) RSG_TRY(tile_build_split(h, p, v.cape, strict_split));
    pipe_maxlen_kernel<<<(p.nblk * 32 + 255) / 256, 256, 0, st>>>(p.nblk, p.base, p.nseg > 0 ? v.cape : 0, p.rowptr, p.pdesc);
    pipe_rowptr_copy_kernel<<<(p.rp0_len + 255) / 256, 256, 0, st>>>(m, p.rp0_len, p.base, p.rowptr, p.rp0);
    h->launches += 3;
    if (h->opts[RSG_OPT_SPMV_HALO_START] > 0) {
        pipe_halo_mark_kernel<<<(p.nblk * 32 + 255) / 256, 256, 0, st>>>(p.nblk, p.base, h->opts[RSG_OPT_SPMV_HALO_START], p.colidx, p.pdesc);
        h->launches++;
    }
    RSG_CUDA_TRY(cudaGetLastError());
    return RSG_OK;
}

// resident CTAs per SM the shared-memory footprint allows (what __launch_bounds__ states), at least 42 registers each
template <typename TV, int THREADS, int CAPE, int ROWS> struct TileResidency {
    static constexpr int by_smem = (int)((size_t)(227 * 1024) / (PipeStage<TV, CAPE, ROWS>::bytes + 1024));
    static constexpr int by_regs = 1536 / THREADS;
    static constexpr int value = by_smem < 1 ? 1 : (by_smem < by_regs ? by_smem : by_regs);
};

template <typename TV, typename TX, int THREADS, int CAPE, int ROWS, bool HINT, bool COH_OK>
static int tile_launch_one(rsg_context *h, const SpmvPlan *p, const TV *values, TX alpha, const TX *x, TX beta, TX *y)
{
    const size_t smem = PipeStage<TV, CAPE, ROWS>::bytes;
    constexpr int MINB = HINT ? TileResidency<TV, THREADS, CAPE, ROWS>::value : 0; // 0 = no residency stated (round 1)
    const bool multi = h->wait_flags != nullptr || h->fuse.ctas > 0;
    if (multi && !COH_OK) { h->wait_flags = nullptr; h->fuse.ctas = 0; return RSG_UNSUPPORTED; }
    auto kern = spmv_tile_kernel<TV, TX, THREADS, CAPE, ROWS, MINB, false>;
    auto kern_coh = spmv_tile_kernel<TV, TX, THREADS, CAPE, ROWS, MINB, COH_OK>;
    static PerDeviceOnce configured; // function attributes are per device (and per instantiation)
    if (configured.need(h->device)) {
        RSG_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        if (COH_OK) RSG_CUDA_TRY(cudaFuncSetAttribute(kern_coh, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    }
    HaloFuse fuse = h->fuse;
    // four boundary entries per thread, issued together: the push is a chain of two dependent loads and a remote store per
    // entry, so with few threads walking many entries it outlasts the whole product (measured: 32 CTAs x 9 entries per
    // thread, one at a time = 18 us for a 14 us product)
    if (fuse.ctas > 0) fuse.ctas = std::max(1, std::min(1024, (fuse.nsend + 4 * THREADS - 1) / (4 * THREADS)));
    LongSplit ls = {p->seg, p->lrow, p->seg_part, p->seg_cnt, p->nseg, nullptr};
    if (h->trace && h->trace_ctas >= p->nblk + fuse.ctas + p->nseg) {
        ls.trace = h->trace + (size_t)(h->trace_parity & 1) * 4 * (size_t)h->trace_ctas;
        h->trace_parity ^= 1;
    }
    const int ropt = h->opts[RSG_OPT_SPMV_REDUCE] | (p->lanes_hint == 2 ? 1 << 16 : 0) |
                     ((p->scattered && h->opts[RSG_OPT_SPMV_X_HINT] != 2) ? 1 << 17 : 0); // X_HINT 2: A/B switch of the product staging
    int pdl = h->opts[RSG_OPT_SPMV_PDL]; // see spmv_window.cu / the header
    if (pdl != 1) {
        cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
        if (cudaStreamIsCapturing(h->stream, &cs) != cudaSuccess || cs != cudaStreamCaptureStatusNone) { (void)cudaGetLastError(); pdl = 1; }
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(p->nblk + fuse.ctas + ls.nseg));
    cfg.blockDim = dim3(THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = h->stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl != 1 ? 1 : 0;
    RSG_CUDA_TRY(cudaLaunchKernelEx(&cfg, multi ? kern_coh : kern, p->base, (const int *)p->rp0, p->colidx, values, x, y, alpha, beta,
                                    (const int4 *)p->pdesc, p->first + p->nnz, p->nblk, ropt, pdl == 2 ? 1 : 0, h->wait_flags,
                                    h->wait_flags ? h->wait_n : 0, h->wait_epoch, h->err_flag, wait_spins(h), fuse, ls));
    h->wait_flags = nullptr; // one-shot
    h->fuse.ctas = 0;
    h->launches++;
    h->last_spmv_blocks = p->nblk;
    return RSG_OK;
}

template <typename TV, typename TX>
int tile_launch(rsg_context *h, const SpmvPlan *p, const TV *values, TX alpha, const TX *x, TX beta, TX *y)
{
    switch (p->variant) { // the shapes AUTO picks (0, 1, 5) also exist as multi-GPU (coherent-gather) instantiations
    case 1: return tile_launch_one<TV, TX, 128, 4352, 512, true, true>(h, p, values, alpha, x, beta, y);
    case 2: return tile_launch_one<TV, TX, 64, 1152, 256, true, false>(h, p, values, alpha, x, beta, y);
    case 3: return tile_launch_one<TV, TX, 256, 4352, 512, true, false>(h, p, values, alpha, x, beta, y);
    case 4: return tile_launch_one<TV, TX, 256, 2304, 512, true, false>(h, p, values, alpha, x, beta, y);
    case 5: return tile_launch_one<TV, TX, 128, 1152, 256, true, true>(h, p, values, alpha, x, beta, y);
    case 6: return tile_launch_one<TV, TX, 128, 2304, 512, false, false>(h, p, values, alpha, x, beta, y);
    case 7: return tile_launch_one<TV, TX, 128, 4352, 512, false, false>(h, p, values, alpha, x, beta, y);
    case 8: return tile_launch_one<TV, TX, 128, 768, 128, true, false>(h, p, values, alpha, x, beta, y);
    case 9: return tile_launch_one<TV, TX, 64, 512, 64, true, false>(h, p, values, alpha, x, beta, y);
    default: return tile_launch_one<TV, TX, 128, 2304, 512, true, true>(h, p, values, alpha, x, beta, y);
    }
}

template int tile_launch<double, double>(rsg_context *, const SpmvPlan *, const double *, double, const double *, double, double *);
template int tile_launch<float, float>(rsg_context *, const SpmvPlan *, const float *, float, const float *, float, float *);
template int tile_launch<float, double>(rsg_context *, const SpmvPlan *, const float *, double, const double *, double, double *);

} // namespace rsg

// ------------------------------------------------------------------ allocation-free SpMV (spmv_csr_d_noalloc)
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (stream capture, tight serving loops, harnesses that forbid it).  The product is the tile kernel above (shape 0, fp64);
// its plan -- the per-tile descriptors and the zero-based padded row pointers, which the other entry points keep in
// cudaMalloc'ed memory inside the handle -- lives in ONE file-scope device array of fixed size, rebuilt by three small
// kernels on every call (the array belongs to no matrix).  Every GPU that loads the library holds the array until exit.
// Calls that overlap on one GPU (two streams, two threads) race on it; calls on one stream are ordered and safe.
namespace rsg {

constexpr int kNoallocMaxRows = RSG_NOALLOC_MAX_ROWS;   // 1 << 20
constexpr int kNoallocMaxTiles = RSG_NOALLOC_MAX_TILES; // 1 << 16
constexpr size_t kNoallocDescBytes = sizeof(int4) * ((size_t)kNoallocMaxTiles + 1 + 15);       // 256-byte multiple
constexpr size_t kNoallocRpBytes = (sizeof(int) * ((((size_t)kNoallocMaxRows + 1 + 3) & ~(size_t)3) + 4) + 255) & ~(size_t)255;
constexpr size_t kNoallocBytes = kNoallocDescBytes + kNoallocRpBytes;
static_assert(kNoallocDescBytes % 256 == 0 && kNoallocBytes % 256 == 0, "scratch sections stay 256-byte aligned");
static_assert(kNoallocBytes == RSG_NOALLOC_SCRATCH_BYTES, "the header states the size");

// [ tile descriptors | row pointers ]: 1 MB + 4 MB.  In HOST code this name is only a stand-in for the symbol: the device
// address comes from cudaGetSymbolAddress (noalloc_scratch_ptr), once per device.
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoallocBytes];

static unsigned char *g_noalloc_ptr[64]; // device address of the array on each device (index = device ordinal & 63)
static std::atomic<unsigned long long> g_noalloc_ready{0};

static int noalloc_scratch_ptr(int device, unsigned char **out)
{
    const unsigned long long bit = 1ull << (device & 63);
    if (!(g_noalloc_ready.load(std::memory_order_acquire) & bit)) {
        // set-up, with `device` current: the symbol's address on THIS device and the kernel's shared-memory opt-in
        void *ptr = nullptr;
        RSG_CUDA_TRY(cudaGetSymbolAddress(&ptr, g_noalloc_scratch));
        if (!ptr) return RSG_CUDA_ERROR;
        auto kern = spmv_tile_kernel<double, double, 128, 2304, 512, TileResidency<double, 128, 2304, 512>::value, false>;
        RSG_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)PipeStage<double, 2304, 512>::bytes));
        g_noalloc_ptr[device & 63] = static_cast<unsigned char *>(ptr);
        g_noalloc_ready.fetch_or(bit, std::memory_order_release);
    }
    *out = g_noalloc_ptr[device & 63];
    return RSG_OK;
}

} // namespace rsg

extern "C" int spmv_csr_d_noalloc(int m, int n, int nnz, int base, const int *rowptr, const int *colidx, const double *values,
                                  double alpha, const double *x, double beta, double *y, void *cuda_stream)
{
    using namespace rsg;
    if (m < 0 || n < 0 || nnz < 0 || (base != 0 && base != 1) || !rowptr || (m > 0 && !y) || (nnz > 0 && (!colidx || !values || !x)))
        return RSG_BAD_ARG;
    if ((((uintptr_t)colidx) & 15) != 0 || (((uintptr_t)values) & 15) != 0) return RSG_BAD_ARG; // the bulk copies need 16-byte aligned arrays
    if (m > kNoallocMaxRows) return RSG_SIZE_LIMIT;
    if (m == 0) return RSG_OK;
    // the tile size of pipe_build_plan for shape 0 under default options (nnz is trusted here: no device read-back)
    constexpr int THREADS = 128, CAPE = 2304, ROWS = 512;
    const int tmax = CAPE * 8 / 9;
    const int w = (tmax + ROWS - 1) / ROWS;
    const double avg = (double)nnz / m;
    int lanes = 1;
    if (avg >= 12.0) {
        const double rows_fit = tmax / (avg + w);
        if (THREADS / rows_fit >= 1.5 && THREADS / rows_fit < 3.0) lanes = 2;
    }
    const long long want = (long long)((avg + w) * (THREADS / lanes) + 0.5);
    const int tile = (int)std::min<long long>(tmax, std::max<long long>(want, 512));
    const long long items = (long long)w * m + nnz;
    const long long ntiles_ll = std::max<long long>(1, (items + tile - 1) / tile);
    if (ntiles_ll > kNoallocMaxTiles) return RSG_SIZE_LIMIT;
    const int ntiles = (int)ntiles_ll;
    const int rp0_len = ((m + 1 + 3) & ~3) + 4;

    int device = 0;
    RSG_CUDA_TRY(cudaGetDevice(&device));
    unsigned char *scratch = nullptr;
    RSG_TRY(noalloc_scratch_ptr(device, &scratch));
    int4 *desc = reinterpret_cast<int4 *>(scratch);
    int *rp0 = reinterpret_cast<int *>(scratch + kNoallocDescBytes);
    cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
    // every word the product reads is rewritten here (descriptors 0..ntiles incl. .z/.w, row pointers 0..rp0_len-1):
    // nothing of the previous call's contents survives into this one
    pipe_partition_kernel<<<(ntiles + 1 + 255) / 256, 256, 0, st>>>(m, base, rowptr, tile, w, ntiles, desc);
    pipe_maxlen_kernel<<<(ntiles * 32 + 255) / 256, 256, 0, st>>>(ntiles, base, 0, rowptr, desc);
    pipe_rowptr_copy_kernel<<<(rp0_len + 255) / 256, 256, 0, st>>>(m, rp0_len, base, rowptr, rp0);
    HaloFuse fuse;
    std::memset(&fuse, 0, sizeof(fuse));
    LongSplit ls = {nullptr, nullptr, nullptr, nullptr, 0, nullptr}; // rows longer than a stage are reduced CTA-per-row inside their tile
    const int ropt = lanes == 2 ? 1 << 16 : 0;
    constexpr int MINB = TileResidency<double, THREADS, CAPE, ROWS>::value;
    spmv_tile_kernel<double, double, THREADS, CAPE, ROWS, MINB, false>
        <<<ntiles, THREADS, PipeStage<double, CAPE, ROWS>::bytes, st>>>(base, (const int *)rp0, colidx, values, x, y, alpha, beta,
                                                                          (const int4 *)desc, nnz, ntiles, ropt, 0, nullptr, 0, 0,
                                                                          nullptr, 0, fuse, ls);
    RSG_CUDA_TRY(cudaGetLastError());
    return RSG_OK;
}
