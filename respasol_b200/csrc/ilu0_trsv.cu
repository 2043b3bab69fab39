// ilu0_trsv.cu -- "Numeric" ILU(0) factorization and the L/U triangular solves for sm_100a,
// replacing cusparse?csrilu02 and the two cusparse?csrsv2_solve calls of the reference harness
// (/root/reference/GPU/ilu0.cu:257-310).  Semantics: in place, natural order, no pivoting, zero
// fill; L unit-lower (diagonal not stored), U upper incl. diagonal; the second solve is the
// INTENDED U solve (the reference's call is an L^T solve left over from an IC0 example).
//
// Both phases are driven by the level-set schedule built on device (levels.cu): rows are handed
// out through an atomic ticket in level order, so rows of one level run side by side and their
// dependencies -- all in earlier levels -- have almost always finished; correctness does not
// rest on that: every dependency is still guarded by a per-row flag (factorization) or by the
// solution entry itself carrying a not-yet-written sentinel (solves).  A row only waits on rows
// whose ticket was drawn earlier by a warp that is resident, hence no deadlock and no grid-wide
// barrier per level.  Waiting never spins inside a divergent branch: a warp sweeps all its rows,
// those that are ready advance, then the warp votes and sweeps again.
//
// Arithmetic follows the oracle exactly (k ascending, explicit fma, IEEE division), so the
// factors and the thread-per-row solves are bit-identical to the CPU restatement.
#include "rsg_internal.cuh"

namespace rsg {

// a waiting row gives up after `spin_limit` sweeps (RSG_OPT_WAIT_SPINS, default 2^22: a malformed schedule or input)
// measured on config 3 (tools/tri_sweep.py, profiles/r1_tri_sweep*.log): 16 entries per poll (a triangle of the
// FEM-like matrix has up to 13 entries per row: one L2 round trip covers the whole row; 8 per poll needed two and
// cost 1.39 ms against 1.05 ms), no gate, and only ~32 levels' worth of rows in flight -- every extra waiting row
// is polling traffic in L2 that slows the rows on the dependency chain down (2.8 ms with every row resident).
#define RSG_TRSV_DEFAULT_BATCH 16
#define RSG_TRSV_DEFAULT_GATE 0

// ------------------------------------------------------------------ flags / sentinels
__device__ __forceinline__ int ld_acquire(const int *p)
{
    int v;
    asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release(int *p, int v)
{
    asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// A solution entry that has not been written yet holds a quiet NaN with a payload no
// arithmetic instruction produces (hardware NaN results are canonical).
template <typename T> struct Pending;
template <> struct Pending<double> {
    typedef unsigned long long bits_t;
    static constexpr bits_t kBits = 0xFFF8DEADBEEF0001ULL;
    static __device__ __forceinline__ bits_t load(const double *p)
    {
        bits_t v;
        asm volatile("ld.relaxed.gpu.global.b64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
        return v;
    }
    static __device__ __forceinline__ void store(double *p, double x)
    {
        asm volatile("st.relaxed.gpu.global.b64 [%0], %1;" ::"l"(p), "l"(__double_as_longlong(x)) : "memory");
    }
    static __device__ __forceinline__ double value(bits_t b) { return __longlong_as_double((long long)b); }
};
template <> struct Pending<float> {
    typedef unsigned int bits_t;
    static constexpr bits_t kBits = 0xFFC0DEADu;
    static __device__ __forceinline__ bits_t load(const float *p)
    {
        bits_t v;
        asm volatile("ld.relaxed.gpu.global.b32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
        return v;
    }
    static __device__ __forceinline__ void store(float *p, float x)
    {
        asm volatile("st.relaxed.gpu.global.b32 [%0], %1;" ::"l"(p), "r"(__float_as_uint(x)) : "memory");
    }
    static __device__ __forceinline__ float value(bits_t b) { return __uint_as_float(b); }
};

template <typename T> __device__ __forceinline__ T fma_t(T a, T b, T c);
template <> __device__ __forceinline__ double fma_t(double a, double b, double c) { return fma(a, b, c); }
template <> __device__ __forceinline__ float fma_t(float a, float b, float c) { return fmaf(a, b, c); }
template <typename T> __device__ __forceinline__ T div_t(T a, T b);
template <> __device__ __forceinline__ double div_t(double a, double b) { return __ddiv_rn(a, b); }
template <> __device__ __forceinline__ float div_t(float a, float b) { return __fdiv_rn(a, b); }

// ------------------------------------------------------------------ ILU(0), warp per row
// ctl[0] ticket, ctl[1] timeout flag, ctl[2] smallest row with a zero pivot
//
// Row i needs, for every lower entry k (ascending):  l_ik = a_ik / u_kk,  then  a_ij -= l_ik u_kj
// for the j > k stored in both rows.  Which entry of row k lands on which position of row i is a
// property of the PATTERN, so the warp works that out (the "candidate list": source position in
// `values`, destination slot in the staged row) while the rows it depends on are still being
// factorized -- warps draw tickets several levels ahead of the wavefront.  What is left on the
// dependency chain of a level is: see the flag, fetch the values of the list in ONE round trip,
// run the short fma chain out of shared memory, write the row, release the flag.  Lower entries
// are consumed in ready-prefix runs, so only the entries of the LAST level are processed after
// the final flag arrives.

// Generic path (row or candidate list too long for the staging buffers): everything in global
// memory, one lower entry at a time.
template <typename T>
__device__ __noinline__ void ilu0_row_generic(int lane, int i, int rs, int len, int nl, int base, const int *__restrict__ rowptr,
                                              const int *__restrict__ colidx, T *values, const int *__restrict__ diag_pos,
                                              const int *done, int epoch, int *ctl, int spin_limit, int *err)
{
    const int *rc = colidx + rs;
    volatile T *rv = values + rs;
    for (int c0 = 0; c0 < nl; c0 += 32) {
        // lane <-> lower entry c0+lane: wait for its row, fetch that row's pivot and extent
        const int lk = c0 + lane;
        const bool mine = lk < nl;
        int k = 0, dpk = 0, rek = 0;
        T ukk = 1;
        if (mine) k = rc[lk] - base;
        bool ready = !mine;
        int spins = 0;
        while (!__all_sync(0xffffffffu, ready)) {
            if (!ready) ready = (ld_acquire(done + k) == epoch);
            if (++spins > spin_limit) { if (lane == 0) { atomicExch(ctl + 1, 1); atomicExch(err, 1); } break; }
        }
        __syncwarp();
        if (mine) {
            dpk = diag_pos[k];
            rek = rowptr[k + 1] - base;
            ukk = __ldcg(values + dpk);
        }
        const int cnt = min(32, nl - c0);
        for (int u = 0; u < cnt; ++u) {
            const int dk = __shfl_sync(0xffffffffu, dpk, u);
            const int ek = __shfl_sync(0xffffffffu, rek, u);
            const T pk = __shfl_sync(0xffffffffu, ukk, u);
            T lik = 0;
            if (lane == 0) {
                lik = div_t<T>(rv[c0 + u], pk);
                rv[c0 + u] = lik;
            }
            lik = __shfl_sync(0xffffffffu, lik, 0);
            for (int jj = dk + 1 + lane; jj < ek; jj += 32) {
                const int jraw = colidx[jj];
                const T ukj = __ldcg(values + jj);
                int lo = c0 + u + 1, hi = len;
                while (lo < hi) {
                    const int mid = (lo + hi) >> 1;
                    if (rc[mid] < jraw) lo = mid + 1; else hi = mid;
                }
                if (lo < len && rc[lo] == jraw) rv[lo] = fma_t<T>(-lik, ukj, rv[lo]);
            }
            __syncwarp();
        }
    }
    if (lane == 0 && rv[nl] == (T)0) atomicMin(ctl + 2, i);
}

template <typename T, int WARPS, int WCAP, int LCAP>
__global__ void __launch_bounds__(WARPS * 32)
ilu0_kernel(int n, int base, const int *__restrict__ rowptr, const int *__restrict__ colidx, T *values,
            const int *__restrict__ diag_pos, const int *__restrict__ order, int *done, int epoch, int *ctl, int spin_limit, int *err)
{
    constexpr unsigned short NONE = 0xffff;
    __shared__ int s_col[WARPS][WCAP];            // row i: raw column indices
    __shared__ T s_val[WARPS][WCAP];              // row i: values, updated in place
    __shared__ int s_koff[WARPS][WCAP + 1];       // lower entry u owns candidates [koff[u], koff[u+1])
    __shared__ int s_kb[WARPS][WCAP];             // first upper position of row k_u (diag_pos[k_u] + 1)
    __shared__ int s_src[WARPS][LCAP];            // candidate: position in `values` of u_kj
    __shared__ unsigned short s_dst[WARPS][LCAP]; // candidate: slot of a_ij in the staged row, NONE = no fill
    __shared__ T s_uv[WARPS][LCAP];               // candidate: fetched u_kj
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    int *col = s_col[w], *koff = s_koff[w], *kbeg = s_kb[w], *src = s_src[w];
    unsigned short *dst = s_dst[w];
    volatile T *val = s_val[w];
    T *uv = s_uv[w];
    for (;;) {
        int t = 0;
        if (lane == 0) t = (*(volatile int *)(ctl + 1)) ? n : atomicAdd(ctl, 1); // stop drawing rows after a timeout
        t = __shfl_sync(0xffffffffu, t, 0);
        if (t >= n) break;
        const int i = order[t];
        const int rs = rowptr[i] - base, len = rowptr[i + 1] - base - rs;
        const int nl = diag_pos[i] - rs; // lower entries are [0,nl) of the row (columns sorted)
        bool fast = len <= WCAP;
        int tot = 0;
        if (fast) {
            // ---- structure only: stage the row, size the candidate list
            for (int l = lane; l < len; l += 32) { col[l] = colidx[rs + l]; val[l] = values[rs + l]; }
            __syncwarp();
            for (int c0 = 0; c0 < nl; c0 += 32) {
                const int lk = c0 + lane;
                int cnt = 0, kb = 0;
                if (lk < nl) {
                    const int k = col[lk] - base;
                    kb = diag_pos[k] + 1;
                    cnt = rowptr[k + 1] - base - kb;
                }
                int incl = cnt;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const int v = __shfl_up_sync(0xffffffffu, incl, o);
                    if (lane >= o) incl += v;
                }
                if (lk < nl) { kbeg[lk] = kb; koff[lk] = tot + incl - cnt; }
                tot += __shfl_sync(0xffffffffu, incl, 31);
            }
            if (lane == 0) koff[nl] = tot;
            __syncwarp();
            fast = tot <= LCAP;
        }
        if (!fast) {
            ilu0_row_generic<T>(lane, i, rs, len, nl, base, rowptr, colidx, values, diag_pos, done, epoch, ctl, spin_limit, err);
        } else {
            // ---- structure only: match every upper entry of every row k against row i
            for (int c = lane; c < tot; c += 32) {
                int lo = 0, hi = nl; // u = the lower entry with koff[u] <= c < koff[u+1]
                while (hi - lo > 1) {
                    const int mid = (lo + hi) >> 1;
                    if (koff[mid] <= c) lo = mid; else hi = mid;
                }
                const int sp = kbeg[lo] + (c - koff[lo]);
                const int jraw = colidx[sp];
                int a = lo + 1, b = len;
                while (a < b) {
                    const int mid = (a + b) >> 1;
                    if (col[mid] < jraw) a = mid + 1; else b = mid;
                }
                src[c] = sp;
                dst[c] = (a < len && col[a] == jraw) ? (unsigned short)a : NONE;
            }
            __syncwarp();
            // ---- the dependent part
            bool timed_out = false;
            for (int c0 = 0; c0 < nl && !timed_out; c0 += 32) {
                const int lk = c0 + lane;
                const bool mine = lk < nl;
                const int cnt = min(32, nl - c0);
                const int k = mine ? col[lk] - base : 0;
                bool ready = !mine;
                int processed = 0, spins = 0;
                while (processed < cnt) {
                    if (!ready) ready = (ld_acquire(done + k) == epoch);
                    const unsigned mask = __ballot_sync(0xffffffffu, ready);
                    const unsigned pend = ~(mask >> processed); // bit 0 <-> entry `processed` of the chunk
                    int run = pend ? __ffs((int)pend) - 1 : 32;
                    run = min(run, cnt - processed);
                    if (run == 0) {
                        if (++spins > spin_limit) { if (lane == 0) { atomicExch(ctl + 1, 1); atomicExch(err, 1); } timed_out = true; break; }
                        continue;
                    }
                    __syncwarp(); // the acquiring lanes' observations now order every lane's loads below
                    const int ua = c0 + processed, ub = ua + run;
                    const int cb = koff[ua], ce = koff[ub];
                    for (int c = cb + lane; c < ce; c += 128) {
                        T v[4];
#pragma unroll
                        for (int q = 0; q < 4; ++q) {
                            const int cc = c + 32 * q;
                            v[q] = (cc < ce && dst[cc] != NONE) ? __ldcg(values + src[cc]) : (T)0;
                        }
#pragma unroll
                        for (int q = 0; q < 4; ++q) {
                            const int cc = c + 32 * q;
                            if (cc < ce) uv[cc] = v[q];
                        }
                    }
                    T piv = 1;
                    if (lane >= processed && lane < processed + run) piv = __ldcg(values + kbeg[lk] - 1);
                    __syncwarp();
                    for (int u = ua; u < ub; ++u) {
                        const T pk = __shfl_sync(0xffffffffu, piv, u - c0);
                        const T aik = val[u];
                        __syncwarp();
                        const T lik = div_t<T>(aik, pk);
                        for (int c = koff[u] + lane; c < koff[u + 1]; c += 32) {
                            const unsigned short d = dst[c];
                            if (d != NONE) val[d] = fma_t<T>(-lik, uv[c], val[d]);
                        }
                        if (lane == 0) val[u] = lik;
                        __syncwarp();
                    }
                    processed += run;
                }
            }
            for (int l = lane; l < len; l += 32) values[rs + l] = val[l];
            if (lane == 0 && val[nl] == (T)0) atomicMin(ctl + 2, i);
        }
        // the row is written: publish it (lane 0's release is cumulative over the warp's stores
        // that the warp barrier ordered before it)
        __syncwarp();
        if (lane == 0) st_release(done + i, epoch);
        __syncwarp();
    }
}

// ------------------------------------------------------------------ solves, thread per row
template <typename T> __global__ void fill_pending_kernel(int n, T *a, T *b)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const T s = Pending<T>::value(Pending<T>::kBits);
    a[i] = s;
    if (b) b[i] = s;
}

// UPPER = false:  z_i = b_i - sum_{j<i} l_ij z_j            (unit diagonal)
// UPPER = true :  y_i = (z_i - sum_{j>i} u_ij y_j) / u_ii
// `sol` is pre-filled with the pending sentinel; rhs is b (lower) or z (upper).
// A thread keeps a batch of B entries of its row (column, coefficient) in registers and polls the
// B solution entries in one round trip; the ready prefix is consumed in stored order (the fma
// chain is the oracle's), the rest is polled again on the next sweep without re-reading the row.
// The diagonal of U is fetched before the first wait, off the dependency chain.
// GATE: until the dependency with the deepest level (crit[i], recorded by the level kernel) is
// written, a row polls that ONE entry only -- rows far ahead of the wavefront then cost one L2
// request per sweep instead of B.
template <typename T, bool UPPER, int B, bool GATE>
__global__ void __launch_bounds__(256)
trsv_kernel(int n, int base, const int *__restrict__ rowptr, const int *__restrict__ colidx,
            const T *__restrict__ lu, const int *__restrict__ diag_pos, const int *__restrict__ order,
            const int *__restrict__ crit, const T *__restrict__ rhs, T *sol, int *ctl, int spin_limit, int *err)
{
    typedef Pending<T> P;
    const int lane = threadIdx.x & 31;
    for (;;) {
        int t0 = 0;
        if (lane == 0) t0 = (*(volatile int *)(ctl + 1)) ? n : atomicAdd(ctl, 32);
        t0 = __shfl_sync(0xffffffffu, t0, 0);
        if (t0 >= n) break;
        const int t = t0 + lane;
        const bool have = t < n;
        int i = 0, k = 0, ke = 0, gate = -1;
        T s = 0, dg = 1;
        if (have) {
            i = order[t];
            if (GATE) gate = crit[i];
            const int kd = diag_pos[i];
            if (UPPER) { k = kd + 1; ke = rowptr[i + 1] - base; dg = lu[kd]; }
            else { k = rowptr[i] - base; ke = kd; }
            s = rhs[i];
        }
        int col[B];
        T a[B];
        int nb = 0, head = 0; // cached entries k .. k+nb, of which [0,head) are consumed
        bool done = !have;
        int spins = 0;
        while (!__all_sync(0xffffffffu, done)) {
            if (GATE && !done && gate >= 0 && P::load(sol + gate) != P::kBits) gate = -1;
            if (!done && gate < 0) {
                for (;;) {
                    if (head == nb) {
                        k += nb;
                        head = 0;
                        nb = min(B, ke - k);
                        if (nb <= 0) {
                            P::store(sol + i, UPPER ? div_t<T>(s, dg) : s);
                            done = true;
                            break;
                        }
#pragma unroll
                        for (int u = 0; u < B; ++u)
                            if (u < nb) { col[u] = colidx[k + u] - base; a[u] = lu[k + u]; }
                    }
                    typename P::bits_t v[B];
#pragma unroll
                    for (int u = 0; u < B; ++u)
                        if (u >= head && u < nb) v[u] = P::load(sol + col[u]);
#pragma unroll
                    for (int u = 0; u < B; ++u)
                        if (u == head && u < nb && v[u] != P::kBits) {
                            s = fma_t<T>(-a[u], P::value(v[u]), s);
                            head = u + 1;
                        }
                    if (head < nb) break; // a dependency is still pending: next sweep
                }
            }
            if (++spins > spin_limit) { atomicExch(ctl + 1, 1); atomicExch(err, 1); break; }
        }
    }
}

// Level-by-level variant (RSG_OPT_TRSV_ALGO = 2): rows [lo,hi) of the schedule belong to levels
// that are all complete on entry, plain loads, one launch per run of levels.  A run of thin
// levels is chewed by ONE CTA with a block barrier between levels instead of a launch each.
template <typename T, bool UPPER>
__device__ __forceinline__ void trsv_row_plain(int i, int base, const int *__restrict__ rowptr,
                                               const int *__restrict__ colidx, const T *__restrict__ lu,
                                               const int *__restrict__ diag_pos, const T *rhs, volatile T *sol)
{
    const int kd = diag_pos[i];
    int k, ke;
    if (UPPER) { k = kd + 1; ke = rowptr[i + 1] - base; }
    else { k = rowptr[i] - base; ke = kd; }
    T s = rhs[i];
    for (; k < ke; ++k) s = fma_t<T>(-lu[k], sol[colidx[k] - base], s);
    sol[i] = UPPER ? div_t<T>(s, lu[kd]) : s;
}

template <typename T, bool UPPER>
__global__ void trsv_level_kernel(int lo, int hi, int base, const int *__restrict__ rowptr,
                                  const int *__restrict__ colidx, const T *__restrict__ lu,
                                  const int *__restrict__ diag_pos, const int *__restrict__ order, const T *rhs, T *sol)
{
    const int t = lo + blockIdx.x * blockDim.x + threadIdx.x;
    if (t < hi) trsv_row_plain<T, UPPER>(order[t], base, rowptr, colidx, lu, diag_pos, rhs, sol);
}

template <typename T, bool UPPER>
__global__ void __launch_bounds__(1024)
trsv_thin_levels_kernel(int l0, int l1, const int *__restrict__ level_ptr, int base, const int *__restrict__ rowptr,
                        const int *__restrict__ colidx, const T *__restrict__ lu, const int *__restrict__ diag_pos,
                        const int *__restrict__ order, const T *rhs, T *sol)
{
    __shared__ int s_rows[1024]; // this level's rows, staged while the previous level drains
    int lo = level_ptr[l0];
    for (int l = l0; l < l1; ++l) {
        const int hi = level_ptr[l + 1];
        const int t = lo + threadIdx.x;
        if (t < hi) s_rows[threadIdx.x] = order[t];
        __syncthreads();
        if (t < hi) trsv_row_plain<T, UPPER>(s_rows[threadIdx.x], base, rowptr, colidx, lu, diag_pos, rhs, sol);
        __threadfence_block();
        __syncthreads();
        lo = hi;
    }
}

// ---- chain solve: ONE CTA walks the levels ------------------------------------------------------------------
// The flag-polling kernel above pays, per level, a store that must reach L2, a poll that must find it and usually a
// second poll for the dependencies that were not there yet: 1.6 us per level on config 3, on 148 SMs that are idle most
// of the time.  When levels are narrow (config 3: 323 rows on average, the 2-D 5-point 1000^2: 500) one SM can take a
// whole level, and then a level boundary is a block barrier (tens of ns) instead of a trip through L2 flags:
//   * rows are taken in level order, a "chunk" of at most kChainThreads rows at a time (a wide level is several chunks
//     without a barrier between them, a new level starts with __syncthreads());
//   * everything a row needs except the solution entries is off the dependency chain: its (row, first entry, end, diagonal)
//     record is fetched two chunks ahead, its first B entries (column, coefficient), right-hand side and pivot one chunk
//     ahead, into registers;
//   * the solution entries are plain loads: inside one CTA a barrier orders them after the stores of the previous level
//     (no flags, no sentinel, no fill kernel); the fma chain runs in stored order, so the result is bit-identical to the
//     oracle's and to the polling kernel's.
static constexpr int kChainThreads = 512;
// AUTO: average rows per level up to which one SM beats 148 polling ones.  Measured (tools/trsv_ab.py, profiles/r2_trsv_ab.log):
// 2-D 5-point 200^2 (100 rows per level) 0.72 vs 1.10 ms; config 3 (323 per level) 2.24 vs 1.05 ms, 1000^2 (500) 9.6 vs 5.6 ms --
// one SM's load pipeline (about 35 GB/s of 32-byte sectors here) saturates long before the level barrier matters.
static constexpr int kChainMaxWidth = 128;

__global__ void chain_meta_kernel(int n, int base, int upper, const int *__restrict__ rowptr, const int *__restrict__ diag_pos,
                                  const int *__restrict__ order, int4 *tmeta)
{
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    const int i = order[t], kd = diag_pos[i];
    tmeta[t] = upper ? make_int4(i, kd + 1, rowptr[i + 1] - base, kd) : make_int4(i, rowptr[i] - base, kd, kd);
}

template <typename T, bool UPPER, int B>
__global__ void __launch_bounds__(kChainThreads, 1)
trsv_chain_kernel(int nchunks, const int2 *__restrict__ chunks, const int4 *__restrict__ tmeta, int base,
                  const int *__restrict__ colidx, const T *__restrict__ lu, const T *__restrict__ rhs, T *sol)
{
    const int tid = threadIdx.x;
    auto meta_of = [&](int c, int &newlevel) -> int4 {
        newlevel = 0;
        if (c >= nchunks) return make_int4(-1, 0, 0, 0);
        const int2 ch = __ldg(chunks + c);
        newlevel = (unsigned)ch.x >> 31;
        const int t = (ch.x & 0x7fffffff) + tid;
        return t < ch.y ? __ldg(tmeta + t) : make_int4(-1, 0, 0, 0);
    };
    int col0[B], col1[B];
    T a0[B], a1[B];
    T rhs0 = 0, rhs1 = 0, dg0 = 1, dg1 = 1;
    int nl0, nl1, nl2;
    int4 m0 = meta_of(0, nl0), m1 = meta_of(1, nl1), m2;
    auto load_row = [&](const int4 &m, int (&col)[B], T (&a)[B], T &r, T &dg) {
        if (m.x < 0) return;
#pragma unroll
        for (int u = 0; u < B; ++u)
            if (m.y + u < m.z) { col[u] = __ldg(colidx + m.y + u) - base; a[u] = __ldg(lu + m.y + u); }
        r = __ldg(rhs + m.x);
        if (UPPER) dg = __ldg(lu + m.w);
    };
    load_row(m0, col0, a0, rhs0, dg0);
    for (int c = 0; c < nchunks; ++c) {
        m2 = meta_of(c + 2, nl2);               // two chunks ahead: the record
        load_row(m1, col1, a1, rhs1, dg1);      // one chunk ahead: entries, right-hand side, pivot
        if (m0.x >= 0) {
            T s = rhs0;
            const int len = m0.z - m0.y;
#pragma unroll
            for (int u = 0; u < B; ++u)
                if (u < len) s = fma_t<T>(-a0[u], sol[col0[u]], s);
            for (int k = m0.y + B; k < m0.z; ++k) s = fma_t<T>(-__ldg(lu + k), sol[__ldg(colidx + k) - base], s); // a row longer than B
            sol[m0.x] = UPPER ? div_t<T>(s, dg0) : s;
        }
        if (nl1) __syncthreads(); // the next chunk opens a new level: its rows read what this level wrote
        m0 = m1; nl0 = nl1; rhs0 = rhs1; dg0 = dg1;
#pragma unroll
        for (int u = 0; u < B; ++u) { col0[u] = col1[u]; a0[u] = a1[u]; }
        m1 = m2; nl1 = nl2;
    }
}

template <typename T>
static int chain_prepare(rsg_context *h, TriPlan *p, int which)
{
    if (p->tmeta[which]) return RSG_OK;
    cudaStream_t st = h->stream;
    const int n = p->n;
    const std::vector<int> &lp = p->level_ptr_host[which];
    std::vector<int2> ch;
    for (int l = 0; l < p->nlevels[which]; ++l)
        for (int t = lp[(size_t)l]; t < lp[(size_t)l + 1]; t += kChainThreads)
            ch.push_back(make_int2((int)((unsigned)t | (t == lp[(size_t)l] ? 0x80000000u : 0u)), std::min(t + kChainThreads, lp[(size_t)l + 1])));
    p->nchunks[which] = (int)ch.size();
    RSG_CUDA_TRY(cudaMalloc(&p->tmeta[which], sizeof(int4) * (size_t)(n > 0 ? n : 1)));
    RSG_CUDA_TRY(cudaMalloc(&p->chunks[which], sizeof(int2) * (ch.size() ? ch.size() : 1)));
    if (!ch.empty()) RSG_CUDA_TRY(cudaMemcpyAsync(p->chunks[which], ch.data(), sizeof(int2) * ch.size(), cudaMemcpyHostToDevice, st));
    chain_meta_kernel<<<(n + 255) / 256, 256, 0, st>>>(n, p->base, which, p->rowptr, p->diag_pos, p->order[which], p->tmeta[which]);
    h->launches++;
    RSG_CUDA_TRY(cudaStreamSynchronize(st)); // (ch is a host temporary)
    p->bytes += sizeof(int4) * (size_t)n + sizeof(int2) * ch.size();
    return RSG_OK;
}

template <typename T>
static int trsv_chain(rsg_context *h, TriPlan *p, const T *lu, const T *b, T *z, T *y)
{
    RSG_TRY(chain_prepare<T>(h, p, 0));
    RSG_TRY(chain_prepare<T>(h, p, 1));
    constexpr int B = sizeof(T) == 8 ? 8 : 12;
    trsv_chain_kernel<T, false, B><<<1, kChainThreads, 0, h->stream>>>(p->nchunks[0], p->chunks[0], p->tmeta[0], p->base, p->colidx, lu, b, z);
    trsv_chain_kernel<T, true, B><<<1, kChainThreads, 0, h->stream>>>(p->nchunks[1], p->chunks[1], p->tmeta[1], p->base, p->colidx, lu, z, y);
    h->launches += 2;
    RSG_CUDA_TRY(cudaGetLastError());
    return RSG_OK;
}

// ------------------------------------------------------------------ host side
template <typename T>
int ilu0_run(rsg_context *h, TriPlan *p, T *values, int *zero_pivot_row)
{
    if (zero_pivot_row) *zero_pivot_row = -1;
    if (p->missing_diag_row >= 0) {
        if (zero_pivot_row) *zero_pivot_row = p->missing_diag_row;
        return RSG_ZERO_PIVOT_STRUCTURAL;
    }
    if (p->n == 0) return RSG_OK;
    RSG_TRY(tri_build_levels(h, p, 0));
    cudaStream_t st = h->stream;
    p->epoch++;
    if (p->epoch == 0x7fffffff) { // flags are compared for equality: restart the stamp
        RSG_CUDA_TRY(cudaMemsetAsync(p->done, 0, sizeof(int) * (size_t)p->n, st));
        p->epoch = 1;
    }
    h->pinned_i[8] = 0; h->pinned_i[9] = 0; h->pinned_i[10] = 0x7fffffff; h->pinned_i[11] = 0;
    RSG_CUDA_TRY(cudaMemcpyAsync(p->ctl, h->pinned_i + 8, 4 * sizeof(int), cudaMemcpyHostToDevice, st));
    constexpr int WARPS = 4, WCAP = 128, LCAP = 512;
    auto kern = ilu0_kernel<T, WARPS, WCAP, LCAP>;
    int per_sm = 0;
    RSG_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, WARPS * 32, 0));
    per_sm = std::max(per_sm, 1);
    if (h->opts[RSG_OPT_TRI_CTAS_PER_SM] > 0) per_sm = std::min(per_sm, h->opts[RSG_OPT_TRI_CTAS_PER_SM]);
    const int grid = std::max(1, std::min(h->sm_count * per_sm, (p->n + WARPS - 1) / WARPS));
    kern<<<grid, WARPS * 32, 0, st>>>(p->n, p->base, p->rowptr, p->colidx, values, p->diag_pos, p->order[0], p->done,
                                      p->epoch, p->ctl, wait_spins(h), h->err_flag);
    h->launches++;
    RSG_CUDA_TRY(cudaGetLastError());
    if (!zero_pivot_row) return RSG_OK;
    RSG_CUDA_TRY(cudaMemcpyAsync(h->pinned_i + 12, p->ctl, 4 * sizeof(int), cudaMemcpyDeviceToHost, st));
    RSG_CUDA_TRY(cudaStreamSynchronize(st));
    if (h->pinned_i[13]) { (void)check_err_flag(h); return RSG_DEPENDENCY_TIMEOUT; }
    if (h->pinned_i[14] != 0x7fffffff) {
        *zero_pivot_row = h->pinned_i[14];
        return RSG_ZERO_PIVOT_NUMERIC;
    }
    return RSG_OK;
}

template <typename T, bool UPPER>
static int trsv_levels_launches(rsg_context *h, TriPlan *p, const T *lu, const T *rhs, T *sol)
{
    const int which = UPPER ? 1 : 0;
    const std::vector<int> &lp = p->level_ptr_host[which];
    const int nl = p->nlevels[which];
    cudaStream_t st = h->stream;
    int l = 0;
    while (l < nl) {
        const int width = lp[l + 1] - lp[l];
        if (width > 1024) {
            trsv_level_kernel<T, UPPER><<<(width + 255) / 256, 256, 0, st>>>(lp[l], lp[l + 1], p->base, p->rowptr, p->colidx,
                                                                             lu, p->diag_pos, p->order[which], rhs, sol);
            h->launches++;
            ++l;
        } else {
            int e = l + 1;
            while (e < nl && lp[e + 1] - lp[e] <= 1024) ++e;
            trsv_thin_levels_kernel<T, UPPER><<<1, 1024, 0, st>>>(l, e, p->level_ptr[which], p->base, p->rowptr, p->colidx, lu,
                                                                  p->diag_pos, p->order[which], rhs, sol);
            h->launches++;
            l = e;
        }
    }
    return RSG_OK;
}

template <typename T>
int trsv_run(rsg_context *h, TriPlan *p, const T *lu, const T *b, T *z, T *y)
{
    if (p->missing_diag_row >= 0) return RSG_ZERO_PIVOT_STRUCTURAL;
    if (p->n == 0) return RSG_OK;
    if (!lu || !b || !z || !y || b == z || z == y || b == y) return RSG_BAD_ARG;
    RSG_TRY(tri_build_levels(h, p, -1));
    cudaStream_t st = h->stream;
    const int n = p->n;
    // RSG_OPT_TRSV_ALGO: 0 = by level width (narrow levels: one CTA walks them; wide levels: flag polling on every SM),
    // 1 = flag polling, 2 = one launch per fat level, 3 = chain
    const int talgo = h->opts[RSG_OPT_TRSV_ALGO];
    const long long avg_width = n / std::max(1, std::min(p->nlevels[0], p->nlevels[1]));
    if (talgo == 3 || (talgo == 0 && avg_width <= kChainMaxWidth)) return trsv_chain<T>(h, p, lu, b, z, y);
    if (talgo == 2) {
        RSG_TRY((trsv_levels_launches<T, false>(h, p, lu, b, z)));
        RSG_TRY((trsv_levels_launches<T, true>(h, p, lu, z, y)));
        RSG_CUDA_TRY(cudaGetLastError());
        return RSG_OK;
    }
    RSG_CUDA_TRY(cudaMemsetAsync(p->ctl, 0, 8 * sizeof(int), st));
    fill_pending_kernel<T><<<(n + 255) / 256, 256, 0, st>>>(n, z, y);
    int per_sm = 0;
    int batch = h->opts[RSG_OPT_TRSV_BATCH];
    if (batch != 1 && batch != 2 && batch != 4 && batch != 8 && batch != 16) batch = RSG_TRSV_DEFAULT_BATCH;
    const bool gate = h->opts[RSG_OPT_TRSV_GATE] ? h->opts[RSG_OPT_TRSV_GATE] == 1 : (RSG_TRSV_DEFAULT_GATE != 0);
    int block = h->opts[RSG_OPT_TRSV_BLOCK];
    if (block < 32 || block > 256 || (block & 31)) block = 64;
    typedef void (*kern_t)(int, int, const int *, const int *, const T *, const int *, const int *, const int *, const T *, T *, int *, int, int *);
    kern_t kl = nullptr, ku = nullptr;
#define RSG_PICK(BB, GG) { kl = trsv_kernel<T, false, BB, GG>; ku = trsv_kernel<T, true, BB, GG>; }
    if (gate) { if (batch == 1) RSG_PICK(1, true) else if (batch == 2) RSG_PICK(2, true) else if (batch == 4) RSG_PICK(4, true) else if (batch == 16) RSG_PICK(16, true) else RSG_PICK(8, true) }
    else { if (batch == 1) RSG_PICK(1, false) else if (batch == 2) RSG_PICK(2, false) else if (batch == 4) RSG_PICK(4, false) else if (batch == 16) RSG_PICK(16, false) else RSG_PICK(8, false) }
#undef RSG_PICK
    RSG_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, ku, block, 0));
    per_sm = std::max(per_sm, 1);
    int grid;
    if (h->opts[RSG_OPT_TRI_CTAS_PER_SM] > 0) {
        grid = h->sm_count * std::min(per_sm, h->opts[RSG_OPT_TRI_CTAS_PER_SM]);
    } else {
        // rows in flight ~ 32 x the average level width, at least one CTA per SM
        const long long width = std::max(1, n / std::max(1, std::min(p->nlevels[0], p->nlevels[1])));
        const long long want = (32 * width + block - 1) / block;
        grid = (int)std::min<long long>((long long)h->sm_count * per_sm, std::max<long long>(h->sm_count, want));
    }
    grid = std::max(1, std::min(grid, (n + block - 1) / block));
    kl<<<grid, block, 0, st>>>(n, p->base, p->rowptr, p->colidx, lu, p->diag_pos, p->order[0], p->crit[0], b, z, p->ctl, wait_spins(h), h->err_flag);
    ku<<<grid, block, 0, st>>>(n, p->base, p->rowptr, p->colidx, lu, p->diag_pos, p->order[1], p->crit[1], z, y, p->ctl + 4, wait_spins(h), h->err_flag);
    h->launches += 3;
    RSG_CUDA_TRY(cudaGetLastError());
    return RSG_OK;
}

template int ilu0_run<double>(rsg_context *, TriPlan *, double *, int *);
template int ilu0_run<float>(rsg_context *, TriPlan *, float *, int *);
template int trsv_run<double>(rsg_context *, TriPlan *, const double *, const double *, double *, double *);
template int trsv_run<float>(rsg_context *, TriPlan *, const float *, const float *, float *, float *);

} // namespace rsg
