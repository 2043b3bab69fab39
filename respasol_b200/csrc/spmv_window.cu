// spmv_window.cu -- the x-window SpMV kernel (RSG_SPMV_WINDOW): replaces the cusparseSpMV call of the
// reference harness (/root/reference/GPU/spmv.cu:174-186) for matrices whose neighbouring rows read
// neighbouring x (stencils, FEM orderings, banded matrices -- every SpMV config of BASELINE.json but
// the power-law one).
//
// What round 1's tile kernel waited on (profiles/r1c_c2_tile_fp64_ncu_full.txt): a CTA lived ~4.2 us, of
// which ~1 us was the bulk copy of its tile and ~2.5 us were two dependent batches of x gathers, each a
// round trip to an L2 that is busy streaming the matrix.  Shared memory sat full-but-idle most of the
// time, so the bytes in flight per SM were a third of what the footprint allows.
//
// Here the gathers leave the critical path.  At plan time (win_build_kernel, one CTA per tile) the
// columns a tile touches are reduced to at most 14 contiguous, 16-byte aligned runs of x ("segments":
// a 27-point stencil tile has 3-9, a 5-point one 3), and every column index is rewritten as a 16-bit
// offset into the concatenation of those runs.  At run time one CTA per tile issues, from eight threads
// at once, 1-D bulk async copies (cp.async.bulk -> UBLKCP) of: the row-pointer slice, the 16-bit
// indices, the values, and the x segments -- all completing on one mbarrier.  After the wait the
// product is shared-memory arithmetic only: thread per row (left to right like the oracle: bit-exact)
// or 4..32 lanes per row.  HBM traffic drops as well: 2 instead of 4 bytes per column index.
//
// Multi-GPU: tiles that read x entries a peer writes (columns >= RSG_OPT_SPMV_HALO_START) wait for the
// peers' arrival flags BEFORE their x segments are copied, so no non-coherent load ever touches
// peer-written memory; the fused push CTAs of rsg_spmv_halo_fuse work as in the tile kernel.
#include "spmv_device.cuh"

namespace rsg {

static constexpr int kWinMaxSeg = 14;      // segments per tile record
static constexpr int kWinRecInt4 = 8;      // record = 8 x int4 = 128 bytes: header + 14 x (start, len | off << 16)
static constexpr int kWinBuildThreads = 256;
static constexpr int kWinBitWords = 24576; // builder bitmap: 96 KB = 786432 granules of 4 columns

// Shared-memory layout of one tile (bytes).  Entries are staged from sb = k0 & ~7 (16-byte aligned 16-bit
// indices), nst = staged entry slots; the x window has xtma copied elements + up to 3 tail elements.
struct WinLayout { int nst, o_x, o_idx, o_rp, total; };
__host__ __device__ __forceinline__ WinLayout win_layout(int k0, int nk, int nr, int xtma, int vbytes, int xbytes)
{
    WinLayout L;
    L.nst = ((k0 & 7) + nk + 7) & ~7;
    L.o_x = L.nst * vbytes;                    // multiple of 32
    L.o_idx = L.o_x + (xtma + 4) * xbytes;     // xtma is a multiple of 4
    L.o_rp = L.o_idx + L.nst * 2;              // multiple of 16
    L.total = L.o_rp + ((nr + 1 + 3 + 3) & ~3) * 4;
    return L;
}

struct WinArgs {
    const int4 *desc;       // ntiles records
    const int *rp0;         // zero-based padded row pointers (absolute entry offsets)
    const uint16_t *widx;   // window-local column of entry k at widx[k - first8]
    int first8, nnz_end, n, ntiles;
    unsigned long long *trace; // debug (rsg_debug_trace): 4 x %globaltimer per CTA -- entry, dependency resolved, data staged, done
};



// -------------------------------------------------------------------------------- row reductions
template <typename TV, typename TX>
__device__ __forceinline__ TX win_row_dot(const TV *s_val, const uint16_t *s_idx, const TX *s_x, int k, int e)
{
    TX s = 0;
    for (; k + 4 <= e; k += 4) {
        const unsigned i0 = s_idx[k], i1 = s_idx[k + 1], i2 = s_idx[k + 2], i3 = s_idx[k + 3];
        const TX x0 = s_x[i0], x1 = s_x[i1], x2 = s_x[i2], x3 = s_x[i3];
        const TX a0 = (TX)s_val[k], a1 = (TX)s_val[k + 1], a2 = (TX)s_val[k + 2], a3 = (TX)s_val[k + 3];
        s += a0 * x0; s += a1 * x1; s += a2 * x2; s += a3 * x3;
    }
    for (; k < e; ++k) s += (TX)s_val[k] * s_x[s_idx[k]];
    return s;
}

template <typename TV, typename TX, int THREADS, int G>
__device__ __forceinline__ void win_rows(const TV *s_val, const uint16_t *s_idx, const TX *s_x, const int *rp, int sb, int nr,
                                         int row_glob, TX *__restrict__ y, TX alpha, TX beta)
{
    const int tid = threadIdx.x;
    if (G == 1) { // thread per row, left to right: the oracle's order
        for (int i = tid; i < nr; i += THREADS) {
            const TX s = win_row_dot<TV, TX>(s_val, s_idx, s_x, rp[i] - sb, rp[i + 1] - sb);
            store_y(y, row_glob + i, s, alpha, beta);
        }
    } else {
        constexpr int GROUPS = THREADS / G;
        const int gid = tid / G, lane = tid % G;
        const int iters = (nr + GROUPS - 1) / GROUPS;
        for (int it = 0; it < iters; ++it) {
            const int i = it * GROUPS + gid;
            TX s = 0;
            if (i < nr) {
                const int e = rp[i + 1] - sb;
                for (int k = rp[i] - sb + lane; k < e; k += G) s += (TX)s_val[k] * s_x[s_idx[k]];
            }
#pragma unroll
            for (int o = G / 2; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
            if (i < nr && lane == 0) store_y(y, row_glob + i, s, alpha, beta);
        }
    }
}

// -------------------------------------------------------------------------------- the kernel
// PK: the instantiation whose halo tiles unpack {value, step} packets (rsg_dist_enable_packets); its eight-packets-in-flight loop
// costs registers the single-GPU product must not pay (53 -> 86 would drop a resident CTA per SM), hence a separate one.
template <typename TV, typename TX, int THREADS, int MINB, bool PK>
__global__ void __launch_bounds__(THREADS, MINB)
spmv_window_kernel(const WinArgs w, const TV *__restrict__ values, const TX *__restrict__ x, TX *__restrict__ y, TX alpha, TX beta,
                   int reduce_opt, int early_values, const int *wait_flags, int wait_n, int wait_epoch, int *err_flag, int max_spins,
                   const HaloFuse fuse, const ChainPush chain, const PacketRecv pk)
{
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ __align__(16) int4 s_desc[kWinRecInt4];
    __shared__ __align__(8) uint64_t s_bar;
    const int tid = threadIdx.x;
    // Programmatic dependent launch: the NEXT kernel of the stream may start its CTAs as soon as every CTA of this grid
    // has got here (they fill the SM slots this grid's last wave frees), and this grid was allowed to start the same way:
    // everything up to grid_dependency_wait() below reads plan-owned, immutable data only (tile record, row pointers,
    // 16-bit columns -- and the values when the caller promised them static), so descriptor latency, the launch gap and
    // the ramp of a product hide behind the tail of the one before it.
    // (pull exchange: the tiles trigger only once their own dependency is resolved -- the next product's fetch CTAs write the
    // halo of the image the product BEFORE this one reads, so they must not start while that one may still be running)
    if (!fuse.pull || (int)blockIdx.x < fuse.ctas) grid_launch_dependents();
    unsigned long long *tr = w.trace ? w.trace + 4 * (size_t)blockIdx.x : nullptr;
    if (tr && tid == 0) tr[0] = global_ns();
    if ((int)blockIdx.x < fuse.ctas) { // fused exchange: publish this rank's boundary entries of x to the peers, or fetch theirs
        if (fuse.pull) {
            halo_pull<TX, THREADS>(fuse, err_flag, max_spins);
            if (tr && tid == 0) tr[1] = tr[2] = tr[3] = global_ns();
            return;
        }
        grid_dependency_wait(); // x is what the work before this launch produced
        if (tr && tid == 0) tr[1] = global_ns();
        halo_fuse_push<TX, THREADS>(fuse, x);
        if (tr && tid == 0) tr[2] = tr[3] = global_ns();
        return;
    }
    const int tile = (int)blockIdx.x - fuse.ctas;
    int4 d = make_int4(0, 0, 0, 0);
    if (tid < kWinRecInt4) { // the 128-byte record: one request
        d = __ldg(w.desc + (size_t)tile * kWinRecInt4 + tid);
        s_desc[tid] = d;
    }
    if (tid == 0) mbar_init(&s_bar, 1);
    __syncthreads();
    const int4 hd = s_desc[0];
    const int r0 = hd.x, k0 = hd.y;
    const int nr = hd.z & 0xffff, nseg = (hd.z >> 16) & 0xff, flags = (hd.z >> 24) & 0xff;
    const int nk = hd.w & 0xffff, xtma = (int)((unsigned)hd.w >> 16);
    const bool must_wait = wait_n > 0 && ((flags & 1) || tile == w.ntiles - 1);
    // packet exchange: the halo entries of this tile's x window are unpacked out of {value, epoch} packets by the CTA's
    // threads (each polls the entries it copies); only the local part of the window is bulk-copied
    const bool pk_tile = PK && pk.pkts != nullptr && (flags & 1);
    if (nr == 0 && !must_wait) return;
    const WinLayout L = win_layout(k0, nk, nr, xtma, (int)sizeof(TV), (int)sizeof(TX));
    TV *s_val = reinterpret_cast<TV *>(smem);
    TX *s_x = reinterpret_cast<TX *>(smem + L.o_x);
    uint16_t *s_idx = reinterpret_cast<uint16_t *>(smem + L.o_idx);
    const int sb = k0 & ~7, k1 = k0 + nk;
    const int qe = max(sb, min(sb + L.nst, w.nnz_end & ~7)); // whole octets only, never past the end of the arrays
    const uint32_t nstage = (uint32_t)(qe - sb);
    const int ra = r0 & ~3, rp_shift = r0 - ra;
    const uint32_t nrp = (uint32_t)((rp_shift + nr + 1 + 3) & ~3);
    if (tid == 0 && nr > 0) {
        const uint64_t spol = make_evict_first_policy();
        int xcopied = xtma;
        if (pk_tile) { // what the bulk copies will NOT bring: the part of every segment at or beyond halo_start
            for (int j = 0; j < nseg; ++j) {
                const int4 q = s_desc[1 + (j >> 1)];
                const int st = (j & 1) ? q.z : q.x, len = (int)((unsigned)((j & 1) ? q.w : q.y) & 0xffffu);
                xcopied -= max(0, st + len - max(st, pk.halo_start));
            }
        }
        mbar_expect_tx(&s_bar, nrp * 4u + nstage * (uint32_t)(sizeof(TV) + 2) + (uint32_t)xcopied * (uint32_t)sizeof(TX));
        bulk_g2s(smem + L.o_rp, w.rp0 + ra, nrp * 4u, &s_bar, spol);
        if (nstage) {
            bulk_g2s(s_idx, w.widx + (sb - w.first8), nstage * 2u, &s_bar, spol);
            if (early_values) bulk_g2s(s_val, values + sb, nstage * (uint32_t)sizeof(TV), &s_bar, spol);
        }
    }
    grid_dependency_wait(); // from here on: x, y and (unless promised static) the values -- what earlier work may have written
    if (fuse.pull) grid_launch_dependents();
    if (tr && tid == 0) tr[1] = global_ns();
    if (tid == 0 && nr > 0 && nstage && !early_values)
        bulk_g2s(s_val, values + sb, nstage * (uint32_t)sizeof(TV), &s_bar, make_evict_first_policy());
    if (must_wait) {
        // x entries of this tile are written by peers over NVLink: wait for their arrival flags (the matrix part of
        // the tile is already in flight), THEN copy the segments -- the copy is the first read of that memory
        halo_flag_wait(wait_flags, wait_n, wait_epoch, err_flag, max_spins);
        __syncthreads();
        if (nr == 0) return;
    }
    if (tid >= 1 && tid < kWinRecInt4) { // seven threads issue two x segments each
        const uint64_t xpol = make_evict_last_policy();
        unsigned l0 = (unsigned)d.y & 0xffffu, l1 = (unsigned)d.w & 0xffffu;
        if (pk_tile) { // halo_start is a multiple of 32 entries: the cut keeps every copy a multiple of 16 bytes
            l0 = (unsigned)max(0, min((int)l0, pk.halo_start - d.x));
            l1 = (unsigned)max(0, min((int)l1, pk.halo_start - d.z));
        }
        if (l0) bulk_g2s(s_x + ((unsigned)d.y >> 16), x + d.x, l0 * (uint32_t)sizeof(TX), &s_bar, xpol);
        if (l1) bulk_g2s(s_x + ((unsigned)d.w >> 16), x + d.z, l1 * (uint32_t)sizeof(TX), &s_bar, xpol);
    }
    bool wrote = false;
    if (pk_tile) {
        // A poll is a round trip to an L2 that is busy streaming the matrix, so the halo parts of up to FOUR segments are polled
        // together, one packet of each per thread in flight (a stencil tile has three such segments of about a row each: one
        // round instead of three); stragglers -- a packet still on its way -- are re-polled one by one.
        int ha[4] = {0, 0, 0, 0}, hb[4] = {0, 0, 0, 0}, hd[4] = {0, 0, 0, 0}; // [ha, hb) = halo columns, hd = s_x offset of column 0
        int nh = 0;
        for (int j = 0; j <= nseg; ++j) {
            if (j < nseg) {
                const int4 q = s_desc[1 + (j >> 1)];
                const int st = (j & 1) ? q.z : q.x;
                const unsigned lo = (unsigned)((j & 1) ? q.w : q.y);
                const int a = max(st, pk.halo_start), b = st + (int)(lo & 0xffffu);
                if (a < b) {
#pragma unroll
                    for (int u = 0; u < 4; ++u)
                        if (u == nh) { ha[u] = a; hb[u] = b; hd[u] = (int)(lo >> 16) - st; }
                    ++nh;
                }
                if (nh < 4) continue;
            }
            if (nh == 0) continue;
            int longest = 0;
#pragma unroll
            for (int u = 0; u < 4; ++u) longest = max(longest, hb[u] - ha[u]);
            for (int base = tid; base < longest; base += THREADS) {
                TX v[4];
                bool ok[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) ok[u] = ha[u] + base >= hb[u] || packet_try(pk.pkts, ha[u] + base - pk.halo_start, wait_epoch, &v[u]);
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const int c = ha[u] + base;
                    if (c >= hb[u]) continue;
                    if (!ok[u]) v[u] = packet_wait<TX>(pk.pkts, c - pk.halo_start, wait_epoch, err_flag, max_spins);
                    s_x[hd[u] + c] = v[u];
                }
            }
            nh = 0;
#pragma unroll
            for (int u = 0; u < 4; ++u) { ha[u] = 0; hb[u] = 0; }
        }
        wrote = true;
    }
    if (qe < k1) { // last tile of a matrix whose entry count is not a multiple of 8: < 8 entries by plain loads
        const int k = qe + tid;
        if (tid < 8 && k < k1) { s_val[k - sb] = values[k]; s_idx[k - sb] = w.widx[k - w.first8]; }
        wrote = true;
    }
    if (flags & 2) { // the window reaches the last n % 4 entries of x, which no 16-byte copy may touch
        const int j = nseg - 1;
        const int4 q = s_desc[1 + (j >> 1)];
        const unsigned lo = (unsigned)((j & 1) ? q.w : q.y);
        const int n4 = w.n & ~3;
        if (tid < w.n - n4)
            s_x[(lo >> 16) + (lo & 0xffffu) + tid] = (pk_tile && n4 + tid >= pk.halo_start)
                                                         ? packet_wait<TX>(pk.pkts, n4 + tid - pk.halo_start, wait_epoch, err_flag, max_spins)
                                                         : x[n4 + tid];
        wrote = true;
    }
    if (wrote) __syncthreads();
    mbar_wait(&s_bar, 0);
    if (tr && tid == 0) tr[2] = global_ns();
    const int *rp = reinterpret_cast<const int *>(smem + L.o_rp) + rp_shift;
    int g = reduce_opt & 0xff;
    if (g == 0) { // lanes per row from the tile's row lengths
        const int avg = nk / nr;
        // rows of up to 40 entries: ALWAYS a thread per row, so that their sums run left to right like the oracle's
        // whatever the tile they fall into (bit-exact, and independent of how the matrix was cut)
        g = avg <= 40 ? 1 : avg <= 80 ? 4 : avg <= 160 ? 8 : avg <= 320 ? 16 : 32;
        while (g > 1 && g < 32 && nr * g * 2 <= THREADS) g <<= 1; // few long rows: spread them over the CTA
    }
    switch (g) {
    case 1: win_rows<TV, TX, THREADS, 1>(s_val, s_idx, s_x, rp, sb, nr, r0, y, alpha, beta); break;
    case 2: win_rows<TV, TX, THREADS, 2>(s_val, s_idx, s_x, rp, sb, nr, r0, y, alpha, beta); break;
    case 4: win_rows<TV, TX, THREADS, 4>(s_val, s_idx, s_x, rp, sb, nr, r0, y, alpha, beta); break;
    case 8: win_rows<TV, TX, THREADS, 8>(s_val, s_idx, s_x, rp, sb, nr, r0, y, alpha, beta); break;
    case 16: win_rows<TV, TX, THREADS, 16>(s_val, s_idx, s_x, rp, sb, nr, r0, y, alpha, beta); break;
    default: win_rows<TV, TX, THREADS, 32>(s_val, s_idx, s_x, rp, sb, nr, r0, y, alpha, beta); break;
    }
    if (tr && tid == 0) tr[3] = global_ns() | ((unsigned long long)(flags & 1) << 63);
    if (chain.npeers > 0 && (flags & 4)) {
        // chained products: this tile owns rows a neighbour reads -- their y goes straight into the neighbours' halo slots of
        // the next x image; the last producer tile to finish releases the step number to the neighbours' flags
        __shared__ int s_last;
        __syncthreads(); // the rows above were stored by other threads of this CTA
        for (int i = tid; i < nr; i += THREADS) {
            const int pos = chain.send_pos[r0 + i];
            if (pos < 0) continue;
            const TX v = y[r0 + i];
#pragma unroll 1
            for (int q = 0; q < chain.npeers; ++q) {
                TX *dq = static_cast<TX *>(chain.dst[q]);
                if (dq && pos >= chain.lo[q] && pos < chain.hi[q]) dq[pos] = v;
            }
        }
        __threadfence_system();
        __syncthreads();
        if (tid == 0) {
            const int done = atomicAdd(chain.counter, 1);
            s_last = (done == chain.nprod - 1);
            if (s_last) *chain.counter = 0;
        }
        __syncthreads();
        if (s_last && tid < chain.npeers && chain.flag[tid]) {
            __threadfence_system();
            asm volatile("st.release.sys.global.s32 [%0], %1;" ::"l"(chain.flag[tid]), "r"(chain.epoch) : "memory");
        }
    }
}

// -------------------------------------------------------------------------------- plan: tile cuts
// part[b] = (first row, its zero-based absolute entry offset) of tile b; a tile holds at most `budget` bytes of
// coordinate = entries * eb + rows * rw (plus its last row)
__global__ void win_partition_kernel(int m, int base, const int *__restrict__ rowptr, long long budget, int eb, int rw, int ntiles,
                                     int2 *part)
{
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b > ntiles) return;
    int r = m;
    if (b < ntiles) {
        const int rp0 = rowptr[0];
        const long long target = (long long)b * budget;
        int lo = 0, hi = m;
        while (lo < hi) {
            const int mid = lo + ((hi - lo) >> 1);
            if ((long long)(rowptr[mid] - rp0) * eb + (long long)rw * mid < target) lo = mid + 1; else hi = mid;
        }
        r = lo;
    }
    part[b] = make_int2(r, rowptr[r] - base);
}

// any bit set in positions [lo, hi] of the bitmap (clamped to [0, span))
__device__ __forceinline__ bool win_any_set(const unsigned *bm, int lo, int hi, int span)
{
    lo = max(lo, 0); hi = min(hi, span - 1);
    if (lo > hi) return false;
    const int w0 = lo >> 5, w1 = hi >> 5;
    for (int w = w0; w <= w1; ++w) {
        unsigned mask = 0xffffffffu;
        if (w == w0) mask &= 0xffffffffu << (lo & 31);
        if (w == w1) mask &= 0xffffffffu >> (31 - (hi & 31));
        if (bm[w] & mask) return true;
    }
    return false;
}

// One CTA per tile: the columns of the tile -> bitmap of 4-column granules -> runs (gaps of up to `gap` unused
// granules are bridged, the gap is widened until at most kWinMaxSeg runs remain) -> the 128-byte record and the
// 16-bit window-local index of every entry.  out[0] = max shared memory a tile needs, out[1] = tiles that cannot be
// expressed (column out of range, window too wide or too large), out[2] = tiles that read columns >= halo_start.
__global__ void __launch_bounds__(kWinBuildThreads)
win_build_kernel(int n, int base, const int *__restrict__ colidx, const int2 *__restrict__ part, int first8, int vbytes, int xbytes,
                 int smem_cap, int gap0, int halo_start, const int *__restrict__ send_pos, int4 *wdesc, uint16_t *widx, int *out)
{
    extern __shared__ unsigned bm[];
    __shared__ int s_min, s_max, s_fail, s_halo, s_nstart, s_nend, s_xtma, s_tail, s_prod;
    __shared__ int s_starts[16], s_ends[16], s_off[16], s_len[16];
    constexpr int T = kWinBuildThreads;
    const int tid = threadIdx.x, b = blockIdx.x;
    const int2 p0 = part[b], p1 = part[b + 1];
    const int r0 = p0.x, k0 = p0.y, nr = p1.x - p0.x, nk = p1.y - p0.y;
    if (tid == 0) { s_min = 0x7fffffff; s_max = -1; s_fail = 0; s_halo = 0; s_nstart = 0; s_nend = 0; s_xtma = 0; s_tail = 0; s_prod = 0; }
    __syncthreads();
    if (send_pos) { // does a neighbour read a row of this tile (chained products)?
        int prod = 0;
        for (int i = tid; i < nr; i += T) prod |= send_pos[r0 + i] >= 0;
        if (prod) s_prod = 1;
    }
    {
        int lmin = 0x7fffffff, lmax = -1, bad = 0, halo = 0;
        for (int k = k0 + tid; k < k0 + nk; k += T) {
            const int c = colidx[k] - base;
            if (c < 0 || c >= n) { bad = 1; continue; }
            lmin = min(lmin, c >> 2); lmax = max(lmax, c >> 2);
            if (halo_start > 0 && c >= halo_start) halo = 1;
        }
        if (lmax >= 0) { atomicMin(&s_min, lmin); atomicMax(&s_max, lmax); }
        if (bad) s_fail = 1;
        if (halo) s_halo = 1;
    }
    __syncthreads();
    const int gmin = s_min;
    const int span = nk > 0 && s_max >= 0 ? s_max - gmin + 1 : 0;
    if (tid == 0 && (nk > 65535 || nr > 65535 || span > kWinBitWords * 32)) s_fail = 1;
    __syncthreads();
    if (s_fail) { if (tid == 0) atomicAdd(out + 1, 1); return; }
    int ns = 0;
    if (span > 0) {
        const int words = (span + 31) >> 5;
        for (int wd = tid; wd < words; wd += T) bm[wd] = 0u;
        __syncthreads();
        for (int k = k0 + tid; k < k0 + nk; k += T) {
            const int g = ((colidx[k] - base) >> 2) - gmin;
            atomicOr(&bm[g >> 5], 1u << (g & 31));
        }
        __syncthreads();
        int gap = gap0;
        for (int attempt = 0; attempt < 8; ++attempt) {
            for (int wd = tid; wd < words; wd += T) {
                unsigned bits = bm[wd];
                while (bits) {
                    const int pos = wd * 32 + __ffs(bits) - 1;
                    bits &= bits - 1;
                    if (!win_any_set(bm, pos - 1 - gap, pos - 1, span)) { const int s = atomicAdd(&s_nstart, 1); if (s < 16) s_starts[s] = pos; }
                    if (!win_any_set(bm, pos + 1, pos + 1 + gap, span)) { const int s = atomicAdd(&s_nend, 1); if (s < 16) s_ends[s] = pos; }
                }
            }
            __syncthreads();
            ns = s_nstart;
            if (ns <= kWinMaxSeg) break;
            __syncthreads();
            if (tid == 0) { s_nstart = 0; s_nend = 0; }
            gap = gap * 4 + 3;
            __syncthreads();
        }
        if (ns > kWinMaxSeg) { if (tid == 0) atomicAdd(out + 1, 1); return; }
    }
    if (tid == 0) {
        for (int i = 1; i < ns; ++i) { // insertion sorts (<= 14 items); starts and ends pair up once sorted
            int v = s_starts[i], j = i - 1;
            while (j >= 0 && s_starts[j] > v) { s_starts[j + 1] = s_starts[j]; --j; }
            s_starts[j + 1] = v;
            v = s_ends[i]; j = i - 1;
            while (j >= 0 && s_ends[j] > v) { s_ends[j + 1] = s_ends[j]; --j; }
            s_ends[j + 1] = v;
        }
        int off = 0, tail = 0;
        const int n4 = n & ~3;
        for (int j = 0; j < ns; ++j) {
            const int start_el = (gmin + s_starts[j]) * 4, end_el = (gmin + s_ends[j] + 1) * 4;
            int len = end_el - start_el;
            if (end_el > n) { len = n4 - start_el; tail = 1; } // only the last run can reach the ragged end of x
            s_off[j] = off; s_len[j] = len;
            off += len;
        }
        s_xtma = off; s_tail = tail;
        const WinLayout L = win_layout(k0, nk, nr, off, vbytes, xbytes);
        if (off + 4 > 65535 || L.total > smem_cap) s_fail = 1;
        else atomicMax(out, L.total);
    }
    __syncthreads();
    if (s_fail) { if (tid == 0) atomicAdd(out + 1, 1); return; }
    for (int k = k0 + tid; k < k0 + nk; k += T) {
        const int c = colidx[k] - base;
        const int g = (c >> 2) - gmin;
        int j = 0;
        while (j + 1 < ns && g > s_ends[j]) ++j;
        widx[k - first8] = (uint16_t)(s_off[j] + (c - (gmin + s_starts[j]) * 4));
    }
    int *rec = reinterpret_cast<int *>(wdesc + (size_t)b * kWinRecInt4);
    if (tid == 0) {
        rec[0] = r0; rec[1] = k0;
        rec[2] = nr | (ns << 16) | (((s_halo ? 1 : 0) | (s_tail ? 2 : 0) | (s_prod ? 4 : 0)) << 24);
        if (s_prod) atomicAdd(out + 3, 1);
        rec[3] = (int)((unsigned)nk | ((unsigned)s_xtma << 16));
        if (s_halo) atomicAdd(out + 2, 1);
    }
    if (tid < kWinMaxSeg) {
        const bool on = tid < ns;
        rec[4 + 2 * tid] = on ? (gmin + s_starts[tid]) * 4 : 0;
        rec[5 + 2 * tid] = on ? (s_len[tid] | (s_off[tid] << 16)) : 0;
    }
}

__global__ void win_rowptr_copy_kernel(int m, int padded, int base, const int *__restrict__ rowptr, int *__restrict__ rp0)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= padded) return;
    rp0[i] = rowptr[min(i, m)] - base;
}

// -------------------------------------------------------------------------------- host side
// RSG_UNSUPPORTED: the matrix has no compact x windows (the caller falls back to the tile kernel).
int window_build_plan(rsg_context *h, SpmvPlan &p, int vbytes, int xbytes, bool strict)
{
    cudaStream_t st = h->stream;
    const int m = p.m;
    const int eb = vbytes + 2;
    int target = h->opts[RSG_OPT_SPMV_WIN_BYTES]; // shared memory per CTA aimed at
    int threads = p.variant == 1 ? 64 : p.variant == 2 ? 256 : 128;
    if (target <= 0) {
        target = 36 * 1024; // 6 CTAs per SM (tools/spmv_sweep.py: 28-45 KB within 1 % on fp64, best for fp32 / mixed)
        // A product of only a few waves of such tiles (configs 1 and 3, a rank's share of config 5 on 8 GPUs) is ramp- and
        // tail-bound: finer tiles with 64-thread CTAs (profiles/r2_window_pdl_sweep.jsonl: C1 13.4 -> 12.8 us, C3 8.5 -> 7.7 us)
        const long long tiles36 = ((long long)p.nnz * eb + 4LL * m) / ((long long)target * 64 / 100) + 1;
        if (p.variant == 0 && tiles36 < 6LL * 6 * h->sm_count) { target = 22 * 1024; threads = 64; }
    }
    target = std::min(std::max(target, 8 * 1024), 160 * 1024);
    int pct = h->opts[RSG_OPT_SPMV_WIN_STREAM_PCT]; // share of it given to the entries (the rest: x window, row pointers)
    if (pct <= 0) pct = 64;
    const long long budget = std::max<long long>(1024, (long long)target * pct / 100);
    if ((long long)p.max_row * eb * 2 > budget) return RSG_UNSUPPORTED; // long rows: the tile kernel splits them
    // rows per tile: at most 8 per thread (5-point rows: 4)
    const int rw = (int)std::max<long long>(4, budget / (8LL * threads));
    const long long total = (long long)p.nnz * eb + (long long)rw * m;
    const long long nt = std::max<long long>(1, (total + budget - 1) / budget);
    if (nt > (1LL << 30)) return RSG_UNSUPPORTED;
    p.nblk = (int)nt;
    p.first8 = p.first & ~7;
    p.rp0_len = ((m + 1 + 3) & ~3) + 4;
    p.xbytes = xbytes;
    int2 *part = nullptr;
    const size_t widx_len = (size_t)(p.first - p.first8) + (size_t)p.nnz + 16;
    RSG_CUDA_TRY(cudaMalloc(&part, sizeof(int2) * ((size_t)p.nblk + 1)));
    cudaError_t e = cudaMalloc(&p.wdesc, sizeof(int4) * kWinRecInt4 * (size_t)p.nblk);
    if (e == cudaSuccess) e = cudaMalloc(&p.widx, sizeof(uint16_t) * widx_len);
    if (e == cudaSuccess) e = cudaMalloc(&p.rp0, sizeof(int) * (size_t)p.rp0_len);
    if (e == cudaSuccess) e = cudaMemsetAsync(p.widx, 0, sizeof(uint16_t) * widx_len, st);
    if (e == cudaSuccess) e = cudaMemsetAsync(p.wdesc, 0, sizeof(int4) * kWinRecInt4 * (size_t)p.nblk, st);
    int *out = h->scratch_i + 32;
    if (e == cudaSuccess) e = cudaMemsetAsync(out, 0, 4 * sizeof(int), st);
    if (e != cudaSuccess) { (void)cudaGetLastError(); cudaFree(part); return RSG_ALLOC_FAILED; }
    win_partition_kernel<<<(p.nblk + 1 + 255) / 256, 256, 0, st>>>(m, p.base, p.rowptr, budget, eb, rw, p.nblk, part);
    static PerDeviceOnce configured; // function attributes are per device (and per instantiation)
    const int bsmem = kWinBitWords * 4;
    if (configured.need(h->device)) {
        if (cudaFuncSetAttribute(win_build_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bsmem) != cudaSuccess) {
            (void)cudaGetLastError(); cudaFree(part); return RSG_CUDA_ERROR;
        }
    }
    const int smem_cap = std::min(h->max_smem_optin - 1024, 200 * 1024);
    int gap0 = h->opts[RSG_OPT_SPMV_WIN_GAP];
    if (gap0 <= 0) gap0 = 3;
    win_build_kernel<<<p.nblk, kWinBuildThreads, bsmem, st>>>(p.n, p.base, p.colidx, part, p.first8, vbytes, xbytes, smem_cap, gap0,
                                                              h->opts[RSG_OPT_SPMV_HALO_START], h->plan_send_pos, p.wdesc, p.widx, out);
    win_rowptr_copy_kernel<<<(p.rp0_len + 255) / 256, 256, 0, st>>>(m, p.rp0_len, p.base, p.rowptr, p.rp0);
    h->launches += 3;
    e = cudaMemcpyAsync(h->pinned_i + 32, out, 4 * sizeof(int), cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    cudaFree(part);
    if (e != cudaSuccess) { (void)cudaGetLastError(); return RSG_CUDA_ERROR; }
    if (h->pinned_i[33] > 0) return RSG_UNSUPPORTED;
    if (strict && h->pinned_i[32] > 2 * target) return RSG_UNSUPPORTED; // x windows far wider than the rows that read them: gathers do better
    p.wsmem = (h->pinned_i[32] + 127) & ~127;
    p.whalo = h->pinned_i[34];
    p.wthreads = threads;
    p.wprod = h->plan_send_pos ? h->pinned_i[35] : 0;
    const bool chained = h->plan_send_pos != nullptr;
    if (!chained && p.whalo > 0 && p.whalo < p.nblk) {
        RSG_TRY(window_place_halo(h, p, h->opts[RSG_OPT_SPMV_HALO_PLACE] > 0 ? h->opts[RSG_OPT_SPMV_HALO_PLACE] : 100, 0));
    } else if (chained && p.wprod > 0 && p.wprod < p.nblk) {
        // multi-GPU tile order (records are self-contained, so the order is free; done once, on the host, stable on both sides).
        // Separate products (the push precedes the product): tiles that read a peer's entries go to the END of the grid -- by
        // the time they are dispatched the neighbours' pushes have landed.  Chained products: boundary tiles (halo readers
        // and producers) go FIRST -- what they read arrived during the previous product, and what they produce has the rest
        // of this product to reach the neighbours.
        std::vector<int4> rec((size_t)p.nblk * kWinRecInt4), out_rec((size_t)p.nblk * kWinRecInt4);
        RSG_CUDA_TRY(cudaMemcpy(rec.data(), p.wdesc, sizeof(int4) * rec.size(), cudaMemcpyDeviceToHost));
        size_t at = 0;
        for (int pass = 0; pass < 2; ++pass)
            for (int b = 0; b < p.nblk; ++b) {
                const int fl = (rec[(size_t)b * kWinRecInt4].z >> 24) & 0xff;
                const bool boundary = chained ? (fl & 5) != 0 : (fl & 1) != 0;
                const bool take = chained ? ((pass == 0) == boundary) : ((pass == 1) == boundary);
                if (take) { std::copy_n(rec.begin() + (size_t)b * kWinRecInt4, kWinRecInt4, out_rec.begin() + at); at += kWinRecInt4; }
            }
        RSG_CUDA_TRY(cudaMemcpy(p.wdesc, out_rec.data(), sizeof(int4) * out_rec.size(), cudaMemcpyHostToDevice));
    }
    p.bytes += sizeof(int4) * kWinRecInt4 * (size_t)p.nblk + sizeof(uint16_t) * widx_len + sizeof(int) * (size_t)p.rp0_len;
    return RSG_OK;
}

// Where the tiles that read a peer's entries sit in the grid (records are self-contained, so the order is free; host side,
// once).  They must not come first (their data is still on its way when the product starts: they would hold SM slots idle),
// and they should not be the very last either: a tile that polls for its halo lives longer than an interior tile, and the
// last tiles of a grid are the critical path into the next launch.  start_pct = share of the grid that is interior tiles
// before the first halo tile; from there halo and interior tiles alternate evenly, and the last wave is interior again.
// start_pct >= 100: all halo tiles at the end (the order the arrival-flag exchange was measured with).
int window_place_halo(rsg_context *h, SpmvPlan &p, int start_pct, int end_pct)
{
    if (p.whalo <= 0 || p.whalo >= p.nblk) return RSG_OK;
    std::vector<int4> rec((size_t)p.nblk * kWinRecInt4), out((size_t)p.nblk * kWinRecInt4);
    RSG_CUDA_TRY(cudaMemcpy(rec.data(), p.wdesc, sizeof(int4) * rec.size(), cudaMemcpyDeviceToHost));
    std::vector<int> in, ha;
    for (int b = 0; b < p.nblk; ++b) {
        const int fl = (rec[(size_t)b * kWinRecInt4].z >> 24) & 0xff;
        ((fl & 1) ? ha : in).push_back(b);
    }
    std::sort(in.begin(), in.end(), [&](int a, int b) { return rec[(size_t)a * kWinRecInt4].x < rec[(size_t)b * kWinRecInt4].x; }); // by first row
    std::sort(ha.begin(), ha.end(), [&](int a, int b) { return rec[(size_t)a * kWinRecInt4].x < rec[(size_t)b * kWinRecInt4].x; });
    std::vector<int> order;
    order.reserve((size_t)p.nblk);
    const int ni = (int)in.size(), nh = (int)ha.size();
    if (start_pct >= 100) {
        order = in;
        order.insert(order.end(), ha.begin(), ha.end());
    } else {
        int tail = std::min(ni / 4, 6 * h->sm_count);
        if (h->opts[RSG_OPT_SPMV_HALO_END] > 0) end_pct = h->opts[RSG_OPT_SPMV_HALO_END];
        if (end_pct > 0 && end_pct < 100) tail = std::min(ni, std::max(tail, (int)((long long)p.nblk * (100 - end_pct) / 100)));
        const int head = std::min(ni - tail, (int)((long long)p.nblk * std::max(start_pct, 0) / 100));
        const int mid = ni - tail - head; // interior tiles that alternate with the halo tiles
        for (int i = 0; i < head; ++i) order.push_back(in[(size_t)i]);
        int ii = 0, hh = 0;
        while (ii < mid || hh < nh) { // even merge: after k of the (mid + nh) slots, about k * nh / (mid + nh) halo tiles are out
            const long long k = ii + hh + 1;
            if (hh < nh && (ii >= mid || (long long)hh * (mid + nh) < k * nh)) order.push_back(ha[(size_t)hh++]);
            else order.push_back(in[(size_t)(head + ii++)]);
        }
        for (int i = head + mid; i < ni; ++i) order.push_back(in[(size_t)i]);
    }
    for (int b = 0; b < p.nblk; ++b) std::copy_n(rec.begin() + (size_t)order[(size_t)b] * kWinRecInt4, kWinRecInt4, out.begin() + (size_t)b * kWinRecInt4);
    RSG_CUDA_TRY(cudaMemcpy(p.wdesc, out.data(), sizeof(int4) * out.size(), cudaMemcpyHostToDevice));
    return RSG_OK;
}

template <typename TV, typename TX, int THREADS>
static int window_launch_one(rsg_context *h, const SpmvPlan *p, const TV *values, TX alpha, const TX *x, TX beta, TX *y)
{
    constexpr int MINB = THREADS >= 256 ? 2 : 4;
    auto kern_plain = spmv_window_kernel<TV, TX, THREADS, MINB, false>;
    auto kern_pk = spmv_window_kernel<TV, TX, THREADS, MINB, true>;
    static PerDeviceOnce configured; // function attributes are per device (and per instantiation)
    if (configured.need(h->device)) {
        RSG_CUDA_TRY(cudaFuncSetAttribute(kern_plain, cudaFuncAttributeMaxDynamicSharedMemorySize, std::min(h->max_smem_optin - 1024, 200 * 1024)));
        RSG_CUDA_TRY(cudaFuncSetAttribute(kern_plain, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
        RSG_CUDA_TRY(cudaFuncSetAttribute(kern_pk, cudaFuncAttributeMaxDynamicSharedMemorySize, std::min(h->max_smem_optin - 1024, 200 * 1024)));
        RSG_CUDA_TRY(cudaFuncSetAttribute(kern_pk, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
    }
    HaloFuse fuse = h->fuse;
    // four boundary entries per thread, issued together: the push is a chain of two dependent loads and a remote store per
    // entry, so with few threads walking many entries it outlasts the whole product (measured: 32 CTAs x 9 entries per
    // thread, one at a time = 18 us for a 14 us product)
    if (fuse.ctas > 0) fuse.ctas = std::max(1, std::min(1024, (fuse.nsend + 4 * THREADS - 1) / (4 * THREADS)));
    ChainPush chain = h->chain;
    if (chain.npeers > 0) chain.nprod = p->wprod;
    h->chain.npeers = 0; // one-shot
    const PacketRecv pk = h->packets;
    h->packets.pkts = nullptr; // one-shot
    WinArgs w = {p->wdesc, p->rp0, p->widx, p->first8, p->first + p->nnz, p->n, p->nblk,
                 nullptr};
    if (h->trace && h->trace_ctas >= p->nblk + fuse.ctas) { // two halves, used in turn: consecutive launches overlap
        w.trace = h->trace + (size_t)(h->trace_parity & 1) * 4 * (size_t)h->trace_ctas;
        h->trace_parity ^= 1;
    }
    // RSG_OPT_SPMV_PDL: 0 = programmatic dependent launch, plan data only ahead of the dependency; 1 = plain launch;
    // 2 = the values too (the caller promises that nothing queued before the product writes them)
    int pdl = h->opts[RSG_OPT_SPMV_PDL];
    if (pdl != 1) { // a stream that is being captured keeps plain launches (the e2e graph of rsg_matrix_spmv_host)
        cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
        if (cudaStreamIsCapturing(h->stream, &cs) != cudaSuccess || cs != cudaStreamCaptureStatusNone) { (void)cudaGetLastError(); pdl = 1; }
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(p->nblk + fuse.ctas));
    cfg.blockDim = dim3(THREADS);
    cfg.dynamicSmemBytes = (size_t)p->wsmem;
    cfg.stream = h->stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl != 1 ? 1 : 0;
    RSG_CUDA_TRY(cudaLaunchKernelEx(&cfg, pk.pkts ? kern_pk : kern_plain, w, values, x, y, alpha, beta, h->opts[RSG_OPT_SPMV_REDUCE], pdl == 2 ? 1 : 0, h->wait_flags,
                                    h->wait_flags ? h->wait_n : 0, h->wait_epoch, h->err_flag, wait_spins(h), fuse, chain, pk));
    h->wait_flags = nullptr; // one-shot
    h->fuse.ctas = 0;
    h->launches++;
    h->last_spmv_blocks = p->nblk;
    return RSG_OK;
}

template <typename TV, typename TX>
int window_launch(rsg_context *h, const SpmvPlan *p, const TV *values, TX alpha, const TX *x, TX beta, TX *y)
{
    if (p->xbytes != (int)sizeof(TX)) return RSG_BAD_ARG;
    switch (p->wthreads) {
    case 64: return window_launch_one<TV, TX, 64>(h, p, values, alpha, x, beta, y);
    case 256: return window_launch_one<TV, TX, 256>(h, p, values, alpha, x, beta, y);
    default: return window_launch_one<TV, TX, 128>(h, p, values, alpha, x, beta, y);
    }
}

template int window_launch<double, double>(rsg_context *, const SpmvPlan *, const double *, double, const double *, double, double *);
template int window_launch<float, float>(rsg_context *, const SpmvPlan *, const float *, float, const float *, float, float *);
template int window_launch<float, double>(rsg_context *, const SpmvPlan *, const float *, double, const double *, double, double *);

} // namespace rsg
