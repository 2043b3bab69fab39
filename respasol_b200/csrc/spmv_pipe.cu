// spmv_pipe.cu -- the pipelined SpMV kernel (RSG_SPMV_PIPE): persistent CTAs, each walking its
// tiles through a ring of STAGES shared-memory buffers.  Replaces the cusparseSpMV call of the
// reference harness (/root/reference/GPU/spmv.cu:174-186) for row lengths without extreme skew.
//
// One elected thread is the producer: for the tile S-1 ahead of the one being reduced it issues
// three 1-D bulk async copies (cp.async.bulk -> UBLKCP): the row-pointer slice, colidx and
// values, all completing on that stage's mbarrier.  The bytes in flight per SM are therefore
// (STAGES-1) tiles per CTA at all times, independent of what the consumer threads do; consumers
// wait on the mbarrier, reduce the tile (thread-, sub-warp- or warp-per-row, x gathered in
// row order so that neighbouring rows of banded matrices read neighbouring x), store y, meet at
// one block barrier and hand the buffer back to the producer.  Nothing on the critical path of
// a tile depends on a global load issued by a consumer except the x gathers themselves.
//
// Plan (built on device by pipe_build_plan): a zero-based, padded copy of the row pointers (so
// that any slice can be bulk-copied from a 16-byte aligned address) and one descriptor per tile
// (first row, first entry, longest row).  Tiles that do not fit a stage (a row longer than the
// tile slack) take a synchronous multi-pass path; a row longer than a whole stage is reduced
// CTA-per-row straight from global memory.
#include "spmv_device.cuh"

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

// A tile that is not "simple" (see the kernel): any row lengths, multi-pass when it does not fit one
// stage, CTA-per-row for a row longer than a stage.  Runs on the NC consumer threads (barrier 1).
template <typename TV, typename TX, int NC, int CAPE, int XH, typename Stage, int BARID = 1, bool LEAN = false>
__device__ __forceinline__ void pipe_general_tile(unsigned char *sp, const PipeMeta &mt, int base, const int *__restrict__ colidx,
                                                  const TV *__restrict__ values, const TX *__restrict__ x, TX *__restrict__ y,
                                                  TX alpha, TX beta, int reduce_opt, uint64_t spol, uint64_t xpol,
                                                  uint64_t *s_aux, uint32_t &aux_parity, bool skip_long = false)
{
    const int tid = threadIdx.x;
    TV *s_val = reinterpret_cast<TV *>(sp + Stage::val_off);
    int *s_col = reinterpret_cast<int *>(sp + Stage::col_off);
    const int *s_rp = reinterpret_cast<const int *>(sp + Stage::rp_off) + mt.rp_shift;
    const int nr = mt.nr, r0 = mt.r0;
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
                group_sync<BARID, NC>();
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
                const TX tot = cta_row_sum<TV, TX, NC, 1, true, BARID>(k0, s_rp[cur + 1], colidx, values, x, base,
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
            group_sync<BARID, NC>();
            if (qe > sb) { mbar_wait(s_aux, aux_parity); aux_parity ^= 1; }
        }
        const int nrows = pe - cur;
        int g = reduce_opt & 0xff;
        const bool prefer2 = (reduce_opt >> 16) & 1; // plan: about threads/2 rows per tile
        const int mode = (reduce_opt >> 8) & 0xff; // RSG_OPT_SPMV_X_HINT on the tile kernel: 3 = the older software-prefetching inner loop (A/B)
        if (g == 0) { // lanes per row from the tile's row-length statistics
            const int avg = (k1 - k0) / nrows;
            const int len = max(avg, mt.maxlen >> 3);
            g = len <= 40 ? 1 : len <= 80 ? 4 : len <= 160 ? 8 : len <= 320 ? 16 : 32;
            while (g < 32 && nrows * g * 2 <= NC) g <<= 1; // few rows: spread them over the CTA
            if (g == 2) g = (avg <= 40 && mt.maxlen <= 64) ? 1 : 4;
            if (prefer2 && g == 1 && mt.maxlen <= 64 && nrows * 2 <= NC) g = 2;
        }
        const int *rp = s_rp + cur;
        if (LEAN) {
            switch (g) {
            case 1: // thread per row: unpredicated whole batches (measured: 8 gathers per batch for fp64 x, 16 for fp32;
                    // tools/ab_option.py SPMV_X_HINT 0 3: config 2 fp64 62.7 -> 60.4 us, fp32 48.2 -> 45.0 us)
                if (mode == 3) reduce_rows_lean<TV, TX, NC, 1>(s_val, s_col, rp, sb, nrows, r0 + cur, x, base, y, alpha, beta);
                else reduce_rows_lean2<TV, TX, NC, (sizeof(TX) == 4 ? 16 : 8)>(s_val, s_col, rp, sb, nrows, r0 + cur, x, base, y, alpha, beta);
                break;
            case 2:
                if (mode == 3) reduce_rows_lean<TV, TX, NC, 2>(s_val, s_col, rp, sb, nrows, r0 + cur, x, base, y, alpha, beta);
                else reduce_rows_lean2g<TV, TX, NC, 2, 8>(s_val, s_col, rp, sb, nrows, r0 + cur, x, base, y, alpha, beta);
                break;
            case 4: reduce_rows_lean<TV, TX, NC, 4>(s_val, s_col, rp, sb, nrows, r0 + cur, x, base, y, alpha, beta); break;
            case 8: reduce_rows_lean<TV, TX, NC, 8>(s_val, s_col, rp, sb, nrows, r0 + cur, x, base, y, alpha, beta); break;
            case 16: reduce_rows_lean<TV, TX, NC, 16>(s_val, s_col, rp, sb, nrows, r0 + cur, x, base, y, alpha, beta); break;
            default: reduce_rows_lean<TV, TX, NC, 32>(s_val, s_col, rp, sb, nrows, r0 + cur, x, base, y, alpha, beta); break;
            }
        } else {
        switch (g) {
            case 1: reduce_rows_gather<TV, TX, NC, 1, XH>(s_val, s_col, rp, sb, nrows, r0 + cur, x, base, xpol, y, alpha, beta); break;
            case 2: reduce_rows_gather<TV, TX, NC, 2, XH>(s_val, s_col, rp, sb, nrows, r0 + cur, x, base, xpol, y, alpha, beta); break;
            case 4: reduce_rows_gather<TV, TX, NC, 4, XH>(s_val, s_col, rp, sb, nrows, r0 + cur, x, base, xpol, y, alpha, beta); break;
            case 8: reduce_rows_gather<TV, TX, NC, 8, XH>(s_val, s_col, rp, sb, nrows, r0 + cur, x, base, xpol, y, alpha, beta); break;
            case 16: reduce_rows_gather<TV, TX, NC, 16, XH>(s_val, s_col, rp, sb, nrows, r0 + cur, x, base, xpol, y, alpha, beta); break;
            default: reduce_rows_gather<TV, TX, NC, 32, XH>(s_val, s_col, rp, sb, nrows, r0 + cur, x, base, xpol, y, alpha, beta); break;
            }
        }
        cur = pe;
        if (!staged || cur < nr) group_sync<BARID, NC>(); // the synchronous path reuses the buffers
        staged = false;
    }
}

// NCW consumer warps + 1 producer warp.  Barrier id 1 = the consumer threads only.
template <typename TV, typename TX, int NCW, int CAPE, int ROWS, int STAGES, int XH, int MINB>
__global__ void __launch_bounds__((NCW + 1) * 32, MINB)
spmv_pipe_kernel(int base, const int *__restrict__ rp0, const int *__restrict__ colidx, const TV *__restrict__ values,
                 const TX *__restrict__ x, TX *__restrict__ y, TX alpha, TX beta, const int4 *__restrict__ desc,
                 int ntiles, int nnz_end, int reduce_opt)
{
    typedef PipeStage<TV, CAPE, ROWS> Stage;
    constexpr int NC = NCW * 32; // consumer threads
    extern __shared__ __align__(128) unsigned char smem_raw[];
    __shared__ __align__(8) uint64_t s_full[STAGES];
    __shared__ __align__(8) uint64_t s_empty[STAGES];
    __shared__ __align__(8) uint64_t s_aux;
    __shared__ PipeMeta s_meta[STAGES];
    const int tid = threadIdx.x;
    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < STAGES; ++s) { mbar_init(&s_full[s], 1); mbar_init(&s_empty[s], NCW); }
        mbar_init(&s_aux, 1);
    }
    __syncthreads();
    // a contiguous run of tiles per CTA: consecutive tiles of a banded matrix share x lines (L1)
    const int t_begin = (int)(((long long)ntiles * blockIdx.x) / gridDim.x);
    const int t_end = (int)(((long long)ntiles * (blockIdx.x + 1)) / gridDim.x);
    const uint64_t spol = make_evict_first_policy();

    if (tid >= NC) { // ---------------------------------------------------------------- producer warp
        if (tid != NC || t_begin >= t_end) return;
        int4 d0 = desc[t_begin], d1 = desc[t_begin + 1];
        int it = 0;
        for (int t = t_begin; t < t_end; ++t, ++it) {
            const int s = it % STAGES;
            // the descriptor of the tile after this one is fetched now and used one iteration later
            const int4 d2 = desc[min(t + 2, ntiles)];
            if (it >= STAGES) mbar_wait(&s_empty[s], (uint32_t)(((it / STAGES) - 1) & 1));
            unsigned char *sp = smem_raw + (size_t)s * Stage::bytes;
            PipeMeta mt;
            mt.r0 = d0.x; mt.nr = d1.x - d0.x; mt.k0 = d0.y; mt.k1 = d1.y; mt.maxlen = d0.z; mt.pad = 0;
            const int ra = mt.r0 & ~3;
            mt.rp_shift = mt.r0 - ra;
            const uint32_t nrp = (uint32_t)((mt.rp_shift + mt.nr + 1 + 3) & ~3);
            const int sb = mt.k0 & ~3;
            // whole quads only; the last quad may reach into the next tile's entries (valid memory)
            // but never past the end of the arrays
            int qe = min((mt.k1 + 3) & ~3, nnz_end & ~3);
            const bool fits = (mt.k1 - sb) <= CAPE && qe > sb;
            if (!fits) qe = sb;
            mt.staged_end = fits ? qe : 0;
            s_meta[s] = mt;
            const uint32_t nq = (uint32_t)(qe - sb);
            mbar_expect_tx(&s_full[s], nrp * 4u + nq * (uint32_t)(sizeof(TV) + sizeof(int)));
            bulk_g2s(sp + Stage::rp_off, rp0 + ra, nrp * 4u, &s_full[s], spol);
            if (nq) {
                bulk_g2s(sp + Stage::val_off, values + sb, nq * (uint32_t)sizeof(TV), &s_full[s], spol);
                bulk_g2s(sp + Stage::col_off, colidx + sb, nq * 4u, &s_full[s], spol);
            }
            d0 = d1; d1 = d2;
        }
        return;
    }

    // ------------------------------------------------------------------------------ consumer warps
    uint64_t xpol = 0;
    if (XH) xpol = make_evict_last_policy();
    uint32_t aux_parity = 0;
    int it = 0;
    for (int t = t_begin; t < t_end; ++t, ++it) {
        const int s = it % STAGES;
        mbar_wait(&s_full[s], (uint32_t)((it / STAGES) & 1));
        unsigned char *sp = smem_raw + (size_t)s * Stage::bytes;
        const PipeMeta mc = s_meta[s];
        pipe_general_tile<TV, TX, NC, CAPE, XH, Stage>(sp, mc, base, colidx, values, x, y, alpha, beta, reduce_opt, spol, xpol,
                                                      &s_aux, aux_parity);
        // this warp is done with stage s: one arrival per consumer warp hands it back to the producer
        __syncwarp();
        if ((tid & 31) == 0) mbar_arrive(&s_empty[s]);
    }
}

// One tile per CTA (RSG_SPMV_TILE): the same plan and staging as the pipelined kernel, but the
// latency is hidden by many small co-resident CTAs instead of a stage ring -- no residency hint in
// __launch_bounds__, ptxas keeps the kernel at ~40 registers, 6-7 CTAs of 128 threads fit an SM.
// Critical path of a CTA: descriptor load -> three bulk copies (row pointers, colidx, values) ->
// mbarrier wait -> reduction.  No consumer-issued global load before the x gathers.
template <typename TV, typename TX, int THREADS, int CAPE, int ROWS>
__global__ void __launch_bounds__(THREADS)
spmv_tile_kernel(int base, const int *__restrict__ rp0, const int *__restrict__ colidx, const TV *__restrict__ values,
                 const TX *__restrict__ x, TX *__restrict__ y, TX alpha, TX beta, const int4 *__restrict__ desc,
                 int nnz_end, int reduce_opt, const int *wait_flags, int wait_n, int wait_epoch, const HaloFuse fuse,
                 const LongSplit ls)
{
    typedef PipeStage<TV, CAPE, ROWS> Stage;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    __shared__ __align__(8) uint64_t s_full;
    __shared__ __align__(8) uint64_t s_aux;
    const int tid = threadIdx.x;
    if ((int)blockIdx.x < fuse.ctas) {
        // ---- fused exchange: these CTAs are dispatched first; they publish this rank's boundary entries of x
        // to the peers (plain stores into P2P-mapped memory over NVLink) while the tiles behind them multiply
        __shared__ int s_last;
        for (int j = blockIdx.x * THREADS + tid; j < fuse.nsend; j += fuse.ctas * THREADS) {
            const TX v = x[fuse.send_idx[j]];
#pragma unroll 1
            for (int q = 0; q < fuse.npeers; ++q) {
                TX *d = static_cast<TX *>(fuse.dst[q]);
                if (d) d[j] = v;
            }
        }
        __threadfence_system();
        __syncthreads();
        if (tid == 0) {
            const int done = atomicAdd(fuse.counter, 1);
            s_last = (done == fuse.ctas - 1);
            if (s_last) *fuse.counter = 0;
        }
        __syncthreads();
        if (s_last && tid < fuse.npeers && fuse.flag[tid]) {
            __threadfence_system();
            asm volatile("st.release.sys.global.s32 [%0], %1;" ::"l"(fuse.flag[tid]), "r"(fuse.epoch) : "memory");
        }
        return;
    }
    const int seg_id = (int)blockIdx.x - fuse.ctas;
    if (seg_id < ls.nseg) {
        // ---- one segment of a long row, straight from global memory with 128-bit loads; the last segment of the
        // row to arrive adds the partial sums in segment order and stores y
        const int4 sg = ls.seg[seg_id];
        const TX tot = cta_row_sum<TV, TX, THREADS, 1, true, 0>(sg.y, sg.z, colidx, values, x, base, reinterpret_cast<TX *>(smem_raw),
                                                               make_evict_first_policy());
        if (tid == 0) {
            const int2 lr = ls.lrow[sg.w];
            *reinterpret_cast<volatile TX *>(ls.part + seg_id) = tot; // one 8-byte slot per segment, whatever TX is
            __threadfence();
            if (atomicAdd(ls.cnt + sg.w, 1) == lr.y - 1) {
                __threadfence();
                TX sum = 0;
                for (int i = 0; i < lr.y; ++i) sum += *reinterpret_cast<volatile TX *>(ls.part + lr.x + i);
                store_y(y, sg.x, sum, alpha, beta);
                ls.cnt[sg.w] = 0;
            }
        }
        return;
    }
    const int tile_id = seg_id - ls.nseg;
    const int4 d0 = desc[tile_id], d1 = desc[tile_id + 1];
    if (tid == 0) { mbar_init(&s_full, 1); mbar_init(&s_aux, 1); }
    __syncthreads();
    const uint64_t spol = make_evict_first_policy();
    PipeMeta mt;
    mt.r0 = d0.x; mt.nr = d1.x - d0.x; mt.k0 = d0.y; mt.k1 = d1.y; mt.maxlen = d0.z; mt.pad = 0;
    if (mt.nr <= 0) return;
    const int ra = mt.r0 & ~3;
    mt.rp_shift = mt.r0 - ra;
    const uint32_t nrp = (uint32_t)((mt.rp_shift + mt.nr + 1 + 3) & ~3);
    const int sb = mt.k0 & ~3;
    int qe = min((mt.k1 + 3) & ~3, nnz_end & ~3);
    const bool fits = (mt.k1 - sb) <= CAPE && qe > sb;
    if (!fits) qe = sb;
    mt.staged_end = fits ? qe : 0;
    if (tid == 0) {
        const uint32_t nq = (uint32_t)(qe - sb);
        mbar_expect_tx(&s_full, nrp * 4u + nq * (uint32_t)(sizeof(TV) + sizeof(int)));
        bulk_g2s(smem_raw + Stage::rp_off, rp0 + ra, nrp * 4u, &s_full, spol);
        if (nq) {
            bulk_g2s(smem_raw + Stage::val_off, values + sb, nq * (uint32_t)sizeof(TV), &s_full, spol);
            bulk_g2s(smem_raw + Stage::col_off, colidx + sb, nq * 4u, &s_full, spol);
        }
    }
    if (d0.w != 0 && wait_n > 0) {
        // this tile gathers x entries that peers write over NVLink: wait for their arrival flags
        // (while the bulk copies of the tile are in flight); every other tile never looks at them
        if (tid < wait_n) {
            int spins = 0, v;
            do {
                asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(wait_flags + tid) : "memory");
            } while (v < wait_epoch && ++spins < (1 << 21));
        }
        __syncthreads();
    }
    mbar_wait(&s_full, 0);
    uint32_t aux_parity = 0;
    pipe_general_tile<TV, TX, THREADS, CAPE, 0, Stage, 0, true>(smem_raw, mt, base, colidx, values, x, y, alpha, beta, reduce_opt,
                                                                spol, 0, &s_aux, aux_parity, ls.nseg > 0);
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
struct PipeVariant { int threads, cape, rows, stages; };
static const PipeVariant kPipeVariants[] = { // consumer threads, entries per stage, rows per stage, stages
    {128, 4352, 512, 3}, {64, 2304, 256, 3}, {32, 1152, 128, 3}, {128, 4352, 512, 2},
    {256, 8704, 1024, 2}, {32, 1152, 128, 4}, {64, 2304, 256, 2}, {96, 3456, 384, 3},
};
static const int kNumPipeVariants = (int)(sizeof(kPipeVariants) / sizeof(kPipeVariants[0]));

int pipe_num_variants() { return kNumPipeVariants; }

static const PipeVariant kTileVariants[] = { // threads, entries per CTA, rows per CTA (stages unused)
    {128, 2304, 512, 1}, {128, 4352, 512, 1}, {64, 1152, 256, 1}, {64, 2304, 512, 1},
    {256, 4352, 512, 1}, {96, 3456, 384, 1}, {192, 6528, 768, 1}, {256, 8704, 1024, 1},
    // short rows (gather-latency bound): small stages so that 56-60 warps stay resident per SM
    {256, 2304, 512, 1}, {128, 1152, 256, 1},
};
static const int kNumTileVariants = (int)(sizeof(kTileVariants) / sizeof(kTileVariants[0]));
int tile_num_variants() { return kNumTileVariants; }

static constexpr int kMaxLongRows = 4096;  // rows split into segments per plan
static constexpr int kSegEntries = 2048;   // entries per segment CTA
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
    const PipeVariant &v = p.algo == RSG_SPMV_TILE ? kTileVariants[p.variant] : kPipeVariants[p.variant];
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

template <typename TV, typename TX, int NC, int CAPE, int ROWS, int STAGES>
static int pipe_launch_one(rsg_context *h, const SpmvPlan *p, const TV *values, TX alpha, const TX *x, TX beta, TX *y)
{
    const size_t smem = PipeStage<TV, CAPE, ROWS>::bytes * STAGES;
    constexpr int THREADS = NC + 32; // + the producer warp
    // resident CTAs per SM the shared memory allows (fp64 figure; ptxas sizes its register budget on it)
    constexpr int MINB = (int)std::min<size_t>(8, std::max<size_t>(1, (size_t)(227 * 1024) / (PipeStage<double, CAPE, ROWS>::bytes * STAGES + 1024)));
    auto k0 = spmv_pipe_kernel<TV, TX, NC / 32, CAPE, ROWS, STAGES, 0, MINB>;
    auto k1 = spmv_pipe_kernel<TV, TX, NC / 32, CAPE, ROWS, STAGES, 1, MINB>;
    static int per_sm = 0; // per instantiation
    if (per_sm == 0) {
        RSG_CUDA_TRY(cudaFuncSetAttribute(k0, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        RSG_CUDA_TRY(cudaFuncSetAttribute(k1, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        int n = 0;
        RSG_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, k0, THREADS, smem));
        per_sm = std::max(n, 1);
    }
    int ctas = per_sm;
    if (h->opts[RSG_OPT_SPMV_CTAS_PER_SM] > 0) ctas = std::min(ctas, h->opts[RSG_OPT_SPMV_CTAS_PER_SM]);
    const int grid = std::max(1, std::min(p->nblk, h->sm_count * ctas));
    const int nnz_end = p->first + p->nnz;
    if (h->opts[RSG_OPT_SPMV_X_HINT])
        k1<<<grid, THREADS, smem, h->stream>>>(p->base, p->rp0, p->colidx, values, x, y, alpha, beta, p->pdesc, p->nblk, nnz_end,
                                               h->opts[RSG_OPT_SPMV_REDUCE]);
    else
        k0<<<grid, THREADS, smem, h->stream>>>(p->base, p->rp0, p->colidx, values, x, y, alpha, beta, p->pdesc, p->nblk, nnz_end,
                                               h->opts[RSG_OPT_SPMV_REDUCE]);
    h->launches++;
    h->last_spmv_blocks = grid;
    return RSG_OK;
}

template <typename TV, typename TX>
int pipe_launch(rsg_context *h, const SpmvPlan *p, const TV *values, TX alpha, const TX *x, TX beta, TX *y)
{
    switch (p->variant) {
    case 1: return pipe_launch_one<TV, TX, 64, 2304, 256, 3>(h, p, values, alpha, x, beta, y);
    case 2: return pipe_launch_one<TV, TX, 32, 1152, 128, 3>(h, p, values, alpha, x, beta, y);
    case 3: return pipe_launch_one<TV, TX, 128, 4352, 512, 2>(h, p, values, alpha, x, beta, y);
    case 4: return pipe_launch_one<TV, TX, 256, 8704, 1024, 2>(h, p, values, alpha, x, beta, y);
    case 5: return pipe_launch_one<TV, TX, 32, 1152, 128, 4>(h, p, values, alpha, x, beta, y);
    case 6: return pipe_launch_one<TV, TX, 64, 2304, 256, 2>(h, p, values, alpha, x, beta, y);
    case 7: return pipe_launch_one<TV, TX, 96, 3456, 384, 3>(h, p, values, alpha, x, beta, y);
    default: return pipe_launch_one<TV, TX, 128, 4352, 512, 3>(h, p, values, alpha, x, beta, y);
    }
}

template <typename TV, typename TX, int THREADS, int CAPE, int ROWS>
static int tile_launch_one(rsg_context *h, const SpmvPlan *p, const TV *values, TX alpha, const TX *x, TX beta, TX *y)
{
    const size_t smem = PipeStage<TV, CAPE, ROWS>::bytes;
    auto kern = spmv_tile_kernel<TV, TX, THREADS, CAPE, ROWS>;
    static bool configured = false; // per instantiation
    if (!configured) {
        RSG_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured = true;
    }
    HaloFuse fuse = h->fuse;
    if (fuse.ctas > 0) fuse.ctas = std::max(1, std::min(32, (fuse.nsend + THREADS - 1) / THREADS));
    LongSplit ls = {p->seg, p->lrow, p->seg_part, p->seg_cnt, p->nseg};
    kern<<<p->nblk + fuse.ctas + ls.nseg, THREADS, smem, h->stream>>>(p->base, p->rp0, p->colidx, values, x, y, alpha, beta, p->pdesc,
                                                                      p->first + p->nnz, h->opts[RSG_OPT_SPMV_REDUCE] | (h->opts[RSG_OPT_SPMV_X_HINT] << 8) | (p->lanes_hint == 2 ? 1 << 16 : 0), h->wait_flags,
                                                                      h->wait_flags ? h->wait_n : 0, h->wait_epoch, fuse, ls);
    h->wait_flags = nullptr; // one-shot
    h->fuse.ctas = 0;
    h->launches++;
    h->last_spmv_blocks = p->nblk;
    return RSG_OK;
}

template <typename TV, typename TX>
int tile_launch(rsg_context *h, const SpmvPlan *p, const TV *values, TX alpha, const TX *x, TX beta, TX *y)
{
    switch (p->variant) {
    case 1: return tile_launch_one<TV, TX, 128, 4352, 512>(h, p, values, alpha, x, beta, y);
    case 2: return tile_launch_one<TV, TX, 64, 1152, 256>(h, p, values, alpha, x, beta, y);
    case 3: return tile_launch_one<TV, TX, 64, 2304, 512>(h, p, values, alpha, x, beta, y);
    case 4: return tile_launch_one<TV, TX, 256, 4352, 512>(h, p, values, alpha, x, beta, y);
    case 5: return tile_launch_one<TV, TX, 96, 3456, 384>(h, p, values, alpha, x, beta, y);
    case 6: return tile_launch_one<TV, TX, 192, 6528, 768>(h, p, values, alpha, x, beta, y);
    case 7: return tile_launch_one<TV, TX, 256, 8704, 1024>(h, p, values, alpha, x, beta, y);
    case 8: return tile_launch_one<TV, TX, 256, 2304, 512>(h, p, values, alpha, x, beta, y);
    case 9: return tile_launch_one<TV, TX, 128, 1152, 256>(h, p, values, alpha, x, beta, y);
    default: return tile_launch_one<TV, TX, 128, 2304, 512>(h, p, values, alpha, x, beta, y);
    }
}

template int tile_launch<double, double>(rsg_context *, const SpmvPlan *, const double *, double, const double *, double, double *);
template int tile_launch<float, float>(rsg_context *, const SpmvPlan *, const float *, float, const float *, float, float *);
template int tile_launch<float, double>(rsg_context *, const SpmvPlan *, const float *, double, const double *, double, double *);

template int pipe_launch<double, double>(rsg_context *, const SpmvPlan *, const double *, double, const double *, double, double *);
template int pipe_launch<float, float>(rsg_context *, const SpmvPlan *, const float *, float, const float *, float, float *);
template int pipe_launch<float, double>(rsg_context *, const SpmvPlan *, const float *, double, const double *, double, double *);

} // namespace rsg
