// spmv_device.cuh -- device helpers shared by the SpMV kernels (spmv.cu, spmv_pipe.cu):
// streaming / gather loads with cache hints, mbarrier + 1-D bulk-copy (TMA) wrappers, the
// staging and row-reduction building blocks.
#pragma once

#include "rsg_internal.cuh"

namespace rsg {

// ------------------------------------------------------------------ streaming loads
// HINT 0: ld.global.nc   1: + L1::no_allocate   2: + L2 evict_first policy
template <int HINT> __device__ __forceinline__ int4 ld_stream(const int4 *p, uint64_t pol)
{
    int4 r;
    if (HINT == 0) {
        asm volatile("ld.global.nc.v4.s32 {%0,%1,%2,%3}, [%4];"
                     : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    } else if (HINT == 1) {
        asm volatile("ld.global.nc.L1::no_allocate.v4.s32 {%0,%1,%2,%3}, [%4];"
                     : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    } else {
        asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v4.s32 {%0,%1,%2,%3}, [%4], %5;"
                     : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p), "l"(pol));
    }
    return r;
}
template <int HINT> __device__ __forceinline__ float4 ld_stream(const float4 *p, uint64_t pol)
{
    int4 r = ld_stream<HINT>(reinterpret_cast<const int4 *>(p), pol);
    return make_float4(__int_as_float(r.x), __int_as_float(r.y), __int_as_float(r.z), __int_as_float(r.w));
}
template <int HINT> __device__ __forceinline__ double2 ld_stream(const double2 *p, uint64_t pol)
{
    double2 r;
    if (HINT == 0) {
        asm volatile("ld.global.nc.v2.f64 {%0,%1}, [%2];" : "=d"(r.x), "=d"(r.y) : "l"(p));
    } else if (HINT == 1) {
        asm volatile("ld.global.nc.L1::no_allocate.v2.f64 {%0,%1}, [%2];" : "=d"(r.x), "=d"(r.y) : "l"(p));
    } else {
        asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v2.f64 {%0,%1}, [%2], %3;"
                     : "=d"(r.x), "=d"(r.y) : "l"(p), "l"(pol));
    }
    return r;
}
__device__ __forceinline__ uint64_t make_evict_first_policy()
{
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}

// four consecutive stored values starting at an index that is a multiple of 4
template <int HINT> __device__ __forceinline__ void ld_values4(const double *p, uint64_t pol, double (&v)[4])
{
    double2 a = ld_stream<HINT>(reinterpret_cast<const double2 *>(p), pol);
    double2 b = ld_stream<HINT>(reinterpret_cast<const double2 *>(p) + 1, pol);
    v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y;
}
template <int HINT> __device__ __forceinline__ void ld_values4(const float *p, uint64_t pol, float (&v)[4])
{
    float4 a = ld_stream<HINT>(reinterpret_cast<const float4 *>(p), pol);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w;
}
__device__ __forceinline__ void st_shared4(double *p, const double (&v)[4])
{
    reinterpret_cast<double2 *>(p)[0] = make_double2(v[0], v[1]);
    reinterpret_cast<double2 *>(p)[1] = make_double2(v[2], v[3]);
}
__device__ __forceinline__ void st_shared4(float *p, const float (&v)[4])
{
    *reinterpret_cast<float4 *>(p) = make_float4(v[0], v[1], v[2], v[3]);
}

// ------------------------------------------------------------------ entry staging
// Products a_k * x[col_k] for the entries [k0,k1) go to s_prod[k - (k0 & ~3)].  VEC: whole
// aligned quads use one 128-bit colidx load + 128/256 bits of values per thread, two quads in
// flight per thread; the ragged head and tail quads are done with guarded scalar loads.
template <typename TV, typename TX, int THREADS, int HINT, bool VEC>
__device__ __forceinline__ void stage_products(int k0, int k1, const int *__restrict__ colidx,
                                               const TV *__restrict__ values, const TX *__restrict__ x,
                                               int base, TX *s_prod, uint64_t pol)
{
    const int tid = threadIdx.x;
    if (VEC) {
        const int sb = k0 & ~3;
        const int qf0 = (k0 + 3) >> 2, qf1 = k1 >> 2; // whole quads [qf0,qf1)
        if (qf0 <= qf1) {
            const int he = qf0 << 2, tb = qf1 << 2;
            if (tid < 4) { // head
                const int k = k0 + tid;
                if (k < he) s_prod[k - sb] = (TX)values[k] * __ldg(x + (colidx[k] - base));
            } else if (tid >= 32 && tid < 36) { // tail
                const int k = tb + (tid - 32);
                if (k < k1) s_prod[k - sb] = (TX)values[k] * __ldg(x + (colidx[k] - base));
            }
            for (int q = qf0 + tid; q < qf1; q += 2 * THREADS) {
                const int qb = q + THREADS;
                const bool hb = qb < qf1;
                int4 ca, cb = make_int4(base, base, base, base);
                TV va[4], vb[4] = {0, 0, 0, 0};
                ca = ld_stream<HINT>(reinterpret_cast<const int4 *>(colidx + (q << 2)), pol);
                ld_values4<HINT>(values + (q << 2), pol, va);
                if (hb) {
                    cb = ld_stream<HINT>(reinterpret_cast<const int4 *>(colidx + (qb << 2)), pol);
                    ld_values4<HINT>(values + (qb << 2), pol, vb);
                }
                TX pa[4], pb[4];
                pa[0] = __ldg(x + (ca.x - base)); pa[1] = __ldg(x + (ca.y - base));
                pa[2] = __ldg(x + (ca.z - base)); pa[3] = __ldg(x + (ca.w - base));
                if (hb) {
                    pb[0] = __ldg(x + (cb.x - base)); pb[1] = __ldg(x + (cb.y - base));
                    pb[2] = __ldg(x + (cb.z - base)); pb[3] = __ldg(x + (cb.w - base));
                }
#pragma unroll
                for (int j = 0; j < 4; ++j) pa[j] *= (TX)va[j];
                st_shared4(s_prod + ((q << 2) - sb), pa);
                if (hb) {
#pragma unroll
                    for (int j = 0; j < 4; ++j) pb[j] *= (TX)vb[j];
                    st_shared4(s_prod + ((qb << 2) - sb), pb);
                }
            }
        } else { // [k0,k1) lies inside one quad
            const int k = k0 + tid;
            if (k < k1) s_prod[k - sb] = (TX)values[k] * __ldg(x + (colidx[k] - base));
        }
    } else {
        const int sb = k0 & ~3;
        for (int k = k0 + tid; k < k1; k += THREADS)
            s_prod[k - sb] = (TX)values[k] * __ldg(x + (colidx[k] - base));
    }
}

// ------------------------------------------------------------------ multi-GPU halo helpers (tile + window kernels)
// Coherent x gather for tiles that read entries a peer stores over NVLink while this kernel is resident: a plain
// ld.global (PTX defines ld.global.nc / __ldg only for data that is read-only for the kernel's lifetime).
__device__ __forceinline__ double ld_x_coherent(const double *p)
{
    double r;
    asm volatile("ld.global.f64 %0, [%1];" : "=d"(r) : "l"(p) : "memory");
    return r;
}
__device__ __forceinline__ float ld_x_coherent(const float *p)
{
    float r;
    asm volatile("ld.global.f32 %0, [%1];" : "=f"(r) : "l"(p) : "memory");
    return r;
}

// ---- packet exchange (rsg_dist_enable_packets): an x entry and the step number travel in ONE store, so an entry is usable the
// moment it lands -- no fence, no counter, no flag hop between the data and its consumer (the protocol NCCL calls LL).
// fp64: 16 bytes {lo, epoch, hi, epoch} (both halves checked: 8-byte atomicity suffices); fp32: 8 bytes {bits, epoch}.
__device__ __forceinline__ void packet_store(void *base, int j, double v, int epoch)
{
    const unsigned long long b = (unsigned long long)__double_as_longlong(v);
    uint4 *p = static_cast<uint4 *>(base) + j;
    asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"((unsigned)b), "r"(epoch), "r"((unsigned)(b >> 32)), "r"(epoch) : "memory");
}
__device__ __forceinline__ void packet_store(void *base, int j, float v, int epoch)
{
    uint2 *p = static_cast<uint2 *>(base) + j;
    asm volatile("st.volatile.global.v2.u32 [%0], {%1, %2};" ::"l"(p), "r"(__float_as_uint(v)), "r"(epoch) : "memory");
}
// one poll: true when packet j carries `epoch` (or 0 = a column nobody writes: window padding), value in *v
__device__ __forceinline__ bool packet_try(const void *base, int j, int epoch, double *v)
{
    unsigned a, b, c, d;
    const uint4 *p = static_cast<const uint4 *>(base) + j;
    asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(a), "=r"(b), "=r"(c), "=r"(d) : "l"(p) : "memory");
    *v = __longlong_as_double((long long)(((unsigned long long)c << 32) | a));
    return ((int)b == epoch && (int)d == epoch) || (b == 0u && d == 0u);
}
__device__ __forceinline__ bool packet_try(const void *base, int j, int epoch, float *v)
{
    unsigned a, b;
    const uint2 *p = static_cast<const uint2 *>(base) + j;
    asm volatile("ld.volatile.global.v2.u32 {%0, %1}, [%2];" : "=r"(a), "=r"(b) : "l"(p) : "memory");
    *v = __uint_as_float(a);
    return (int)b == epoch || b == 0u;
}
template <typename TX> __device__ __forceinline__ TX packet_wait(const void *base, int j, int epoch, int *err_flag, int max_spins)
{
    TX v;
    int spins = 0;
    while (!packet_try(base, j, epoch, &v)) {
        if (++spins >= max_spins) { if (err_flag) atomicExch(err_flag, 1); break; }
        if (spins > 64) __nanosleep(spins > 4096 ? 400 : 40);
    }
    return v;
}

// The first `fuse.ctas` CTAs of a product's grid: store this rank's boundary entries of x into the peers' halo slots
// (plain stores into P2P-mapped memory), then the last of them to finish releases `epoch` to the peers' flags.
template <typename TX, int THREADS> __device__ __forceinline__ void halo_fuse_push(const HaloFuse &fuse, const TX *__restrict__ x, int cta = -1)
{
    __shared__ int s_last;
    const int tid = threadIdx.x;
    if (cta < 0) cta = blockIdx.x;
    // four entries per thread, their loads issued back to back: the push is a latency chain, not bandwidth
    const int stride = fuse.ctas * THREADS;
    for (int j0 = cta * THREADS + tid; j0 < fuse.nsend; j0 += 4 * stride) {
        int idx[4];
        TX v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int j = j0 + u * stride;
            idx[u] = -1;
            if (j < fuse.nsend) {
                if (fuse.nruns > 0) { // entry j of the concatenated runs
                    int r = 0, off = j;
                    while (r + 1 < fuse.nruns && off >= fuse.run_len[r]) { off -= fuse.run_len[r]; ++r; }
                    idx[u] = fuse.run_start[r] + off;
                } else {
                    idx[u] = fuse.send_idx[j];
                }
            }
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) v[u] = idx[u] >= 0 ? x[idx[u]] : (TX)0;
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int j = j0 + u * stride;
            if (idx[u] < 0) continue;
#pragma unroll 1
            for (int q = 0; q < fuse.npeers; ++q) {
                TX *d = static_cast<TX *>(fuse.dst[q]);
                if (d && j >= fuse.lo[q] && j < fuse.hi[q]) { // each boundary entry only to the peers that read it
                    if (fuse.packets) packet_store(fuse.dst[q], j, v[u], fuse.epoch); else d[j] = v[u];
                }
            }
        }
    }
    if (fuse.packets) return; // the epoch rode with every entry
    __threadfence_system(); // every thread: its own stores have reached the peers
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
}

template <typename TX> __device__ __forceinline__ TX ld_remote(const TX *p);
template <> __device__ __forceinline__ double ld_remote<double>(const double *p)
{
    double v;
    asm volatile("ld.volatile.global.f64 %0, [%1];" : "=d"(v) : "l"(p) : "memory");
    return v;
}
template <> __device__ __forceinline__ float ld_remote<float>(const float *p)
{
    float v;
    asm volatile("ld.volatile.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
    return v;
}

// Pull exchange (rsg_dist_enable_pull): the first `fuse.ctas` CTAs of a product FETCH this rank's halo.  A remote store has
// to be flushed through NVLink before the grid that issued it counts as complete (measured: the launch dependency of the
// next product resolves 2-4.5 us after the last CTA instead of 1.2 us), and a fence behind remote stores costs 6-10 us while
// the memory system is saturated; a remote LOAD is complete when its data is back.  So: CTA 0 raises this rank's own counter
// ("my x of this step is final") once the launch dependency is resolved; every fetch CTA polls the neighbours' counters over
// NVLink, loads its share of the entries from the owners' x and stores them into this rank's own image; the last one raises
// the local flag the halo tiles wait for.  These CTAs do not wait for this rank's own launch dependency (they write only the
// halo part of an image the product before does not read), so the fetch runs beside the tail of the previous product.
template <typename TX, int THREADS>
__device__ __forceinline__ void halo_pull(const HaloFuse &f, int *err_flag, int max_spins, int cta = -1)
{
    const int tid = threadIdx.x;
    if (cta < 0) cta = blockIdx.x;
    if (cta == 0) {
        asm volatile("griddepcontrol.wait;" ::: "memory");
        if (tid == 0 && f.my_ready) asm volatile("st.release.sys.global.s32 [%0], %1;" ::"l"(f.my_ready), "r"(f.epoch) : "memory");
    }
    if (tid < f.npeers && f.flag[tid]) { // the neighbour's x of this step is final, and it is done with ours of two steps ago
        int spins = 0, v;
        for (;;) {
            asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(f.flag[tid]) : "memory");
            if (v - f.epoch >= 0) break;
            if (++spins >= max_spins) { if (err_flag) atomicExch(err_flag, 1); break; }
            if (spins > 16) __nanosleep(spins > 1024 ? 400 : 100);
        }
    }
    __syncthreads();
    TX *xl = static_cast<TX *>(f.local_x);
    const int stride = f.ctas * THREADS;
    for (int k0 = cta * THREADS + tid; k0 < f.nsend; k0 += 4 * stride) {
        const TX *src[4];
        int col[4];
        TX v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int k = k0 + u * stride;
            src[u] = nullptr;
            col[u] = -1;
            if (k < f.nsend) {
                col[u] = f.pull_col[k];
                const int idx = f.send_idx[k];
#pragma unroll 1
                for (int q = 0; q < f.npeers; ++q)
                    if (f.dst[q] && k >= f.lo[q] && k < f.hi[q]) src[u] = static_cast<const TX *>(f.dst[q]) + idx;
            }
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) v[u] = src[u] ? ld_remote<TX>(src[u]) : (TX)0;
#pragma unroll
        for (int u = 0; u < 4; ++u)
            if (src[u]) xl[col[u]] = v[u];
    }
    __threadfence(); // device scope: the readers are this product's own halo tiles
    __syncthreads();
    if (tid == 0) {
        const int done = atomicAdd(f.counter, 1);
        if (done == f.ctas - 1) {
            *f.counter = 0;
            __threadfence();
            asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(f.local_flag), "r"(f.epoch) : "memory");
        }
    }
}

// Threads tid < n each await flags[tid] >= epoch (acquire, system scope).  A wait that gives up (max_spins polls with
// back-off; the default 2^22 is seconds) raises *err_flag, which the next synchronising call of the handle turns into
// RSG_DEPENDENCY_TIMEOUT.  The caller follows with a block barrier.
__device__ __forceinline__ void halo_flag_wait(const int *flags, int n, int epoch, int *err_flag, int max_spins)
{
    const int tid = threadIdx.x;
    if (tid < n) {
        int spins = 0, v;
        for (;;) {
            asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(flags + tid) : "memory");
            if (v - epoch >= 0) break; // (wrap-aware)
            if (++spins >= max_spins) { if (err_flag) atomicExch(err_flag, 1); break; }
            if (spins > 64) __nanosleep(spins > 4096 ? 400 : 40);
        }
    }
}

__device__ __forceinline__ unsigned long long global_ns() // (rsg_debug_trace)
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// Programmatic dependent launch (PTX griddepcontrol; no-ops in a kernel that was launched without the attribute)
__device__ __forceinline__ void grid_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void grid_dependency_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

// x gather of a row reduction: read-only path, or coherent when peers write x while the kernel runs
template <bool COH, typename TX> __device__ __forceinline__ TX ld_xg(const TX *p) { return COH ? ld_x_coherent(p) : __ldg(p); }

// barrier over the first THREADS threads of the CTA: id 0 = __syncthreads (whole CTA), otherwise a
// named barrier (kernels with a producer warp that does not take part)
template <int BARID, int THREADS> __device__ __forceinline__ void group_sync()
{
    if (BARID == 0) __syncthreads();
    else asm volatile("bar.sync %0, %1;" ::"n"(BARID), "n"(THREADS) : "memory");
}

// CTA-wide sum of a_k*x[col_k] over [k0,k1), result valid in thread 0 (CTA-per-row path).
template <typename TV, typename TX, int THREADS, int HINT, bool VEC, int BARID = 0, bool COH = false>
__device__ __forceinline__ TX cta_row_sum(int k0, int k1, const int *__restrict__ colidx,
                                          const TV *__restrict__ values, const TX *__restrict__ x,
                                          int base, TX *s_red, uint64_t pol)
{
    const int tid = threadIdx.x;
    TX acc = 0;
    int ks = k0, ke = k1;
    if (VEC) {
        const int qf0 = (k0 + 3) >> 2, qf1 = k1 >> 2;
        if (qf0 < qf1) {
            ks = qf0 << 2; ke = qf1 << 2;
            TX acc2 = 0;
            // two quads per thread and iteration, every load of both issued before any is used: a long row's segment is a
            // chain of (index, value) -> x gather -> multiply-add per quad, and with one quad in flight per thread a 2048-entry
            // segment took four dependent round trips (profiles/r2_tile_c4_d_ncu_full.txt: half of the stall samples of config 4)
            for (int q = qf0 + tid; q < qf1; q += 2 * THREADS) {
                const int qb = q + THREADS;
                const bool hb = qb < qf1;
                const int4 ca = ld_stream<HINT>(reinterpret_cast<const int4 *>(colidx + (q << 2)), pol);
                int4 cb = ca;
                TV va[4], vb[4] = {0, 0, 0, 0};
                ld_values4<HINT>(values + (q << 2), pol, va);
                if (hb) {
                    cb = ld_stream<HINT>(reinterpret_cast<const int4 *>(colidx + (qb << 2)), pol);
                    ld_values4<HINT>(values + (qb << 2), pol, vb);
                }
                const TX xa0 = ld_xg<COH>(x + (ca.x - base)), xa1 = ld_xg<COH>(x + (ca.y - base));
                const TX xa2 = ld_xg<COH>(x + (ca.z - base)), xa3 = ld_xg<COH>(x + (ca.w - base));
                const TX xb0 = ld_xg<COH>(x + (cb.x - base)), xb1 = ld_xg<COH>(x + (cb.y - base));
                const TX xb2 = ld_xg<COH>(x + (cb.z - base)), xb3 = ld_xg<COH>(x + (cb.w - base));
                acc += (TX)va[0] * xa0; acc2 += (TX)va[1] * xa1; acc += (TX)va[2] * xa2; acc2 += (TX)va[3] * xa3;
                acc += (TX)vb[0] * xb0; acc2 += (TX)vb[1] * xb1; acc += (TX)vb[2] * xb2; acc2 += (TX)vb[3] * xb3; // (vb = 0 without a second quad)
            }
            acc += acc2;
            // ragged ends
            for (int k = k0 + tid; k < ks; k += THREADS) acc += (TX)values[k] * ld_xg<COH>(x + (colidx[k] - base));
            for (int k = ke + tid; k < k1; k += THREADS) acc += (TX)values[k] * ld_xg<COH>(x + (colidx[k] - base));
        } else {
            for (int k = k0 + tid; k < k1; k += THREADS) acc += (TX)values[k] * ld_xg<COH>(x + (colidx[k] - base));
        }
    } else {
        for (int k = k0 + tid; k < k1; k += THREADS) acc += (TX)values[k] * ld_xg<COH>(x + (colidx[k] - base));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    group_sync<BARID, THREADS>();
    if ((tid & 31) == 0) s_red[tid >> 5] = acc;
    group_sync<BARID, THREADS>();
    TX tot = 0;
    if (tid == 0)
        for (int w = 0; w < THREADS / 32; ++w) tot += s_red[w];
    group_sync<BARID, THREADS>();
    return tot;
}

template <typename TX> __device__ __forceinline__ void store_y(TX *y, int r, TX sum, TX alpha, TX beta)
{
    y[r] = (beta == (TX)0) ? alpha * sum : alpha * sum + beta * y[r];
}

// rows [row_lo,row_lo+nrows) of the tile, entries already staged; G lanes per row
template <typename TX, int THREADS, int G>
__device__ __forceinline__ void reduce_rows(const TX *s_prod, const int *s_rp, int sb, int nrows, int row_glob,
                                            TX *__restrict__ y, TX alpha, TX beta)
{
    const int tid = threadIdx.x;
    if (G == 1) {
        for (int i = tid; i < nrows; i += THREADS) {
            const int a = s_rp[i] - sb, b = s_rp[i + 1] - sb;
            TX s = 0;
            for (int k = a; k < b; ++k) s += s_prod[k];
            store_y(y, row_glob + i, s, alpha, beta);
        }
    } else {
        constexpr int GROUPS = THREADS / G;
        const int gid = tid / G, lane = tid % G;
        const int iters = (nrows + GROUPS - 1) / GROUPS;
        for (int it = 0; it < iters; ++it) {
            const int i = it * GROUPS + gid;
            TX s = 0;
            if (i < nrows) {
                const int a = s_rp[i] - sb, b = s_rp[i + 1] - sb;
                for (int k = a + lane; k < b; k += G) s += s_prod[k];
            }
#pragma unroll
            for (int o = G / 2; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
            if (i < nrows && lane == 0) store_y(y, row_glob + i, s, alpha, beta);
        }
    }
}

// ------------------------------------------------------------------ mbarrier / bulk copy / x gather
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, int count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity)
{
    uint32_t ok;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    } while (!ok);
}
// 1-D bulk copy global -> shared; bytes % 16 == 0, both addresses 16-byte aligned
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, uint32_t bytes, uint64_t *bar, uint64_t pol)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
                 ::"r"(smem_u32(dst)), "l"(__cvta_generic_to_global(src)), "r"(bytes), "r"(smem_u32(bar)), "l"(pol) : "memory");
}
__device__ __forceinline__ uint64_t make_evict_last_policy()
{
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
// "lean" variant of row_dot for kernels that run MANY warps per SM (stream kernel: no residency
// hint in __launch_bounds__, so ptxas keeps it at 32-40 registers and 7 CTAs fit an SM): predicated
// __ldg gathers, a few in flight per lane; the latency is hidden by occupancy instead.  Measured
// faster than the register-hungry batched form whenever the latter costs resident warps.
template <typename TV, typename TX, int G, int B, bool COH = false>
__device__ __forceinline__ TX row_dot_lean(const TV *s_val, const int *s_col, int k, int e, const TX *xb)
{
    TX s = 0;
    TX xa[B];
#pragma unroll
    for (int u = 0; u < B; ++u) xa[u] = (k + u * G < e) ? ld_xg<COH>(xb + s_col[k + u * G]) : (TX)0;
    while (k < e) {
        const int kn = k + B * G;
        TX xn[B];
#pragma unroll
        for (int u = 0; u < B; ++u) xn[u] = (kn + u * G < e) ? ld_xg<COH>(xb + s_col[kn + u * G]) : (TX)0;
#pragma unroll
        for (int u = 0; u < B; ++u)
            if (k + u * G < e) s += (TX)s_val[k + u * G] * xa[u];
#pragma unroll
        for (int u = 0; u < B; ++u) xa[u] = xn[u];
        k = kn;
    }
    return s;
}

// Same sum, same order, fewer instructions: whole batches of B entries run without predicates (B independent
// gathers, then B multiply-adds), only the last partial batch is predicated; the gather address is one IMAD.WIDE
// off an opaque 64-bit base.  (row_dot_lean spends ~17 instructions per entry on a 27-point row: per-entry
// predicates, a 4-instruction 64-bit address chain, and 16 register moves per batch for its software prefetch.)
template <typename TV, typename TX, int G, int B, bool COH = false>
__device__ __forceinline__ TX row_dot_lean2(const TV *s_val, const int *s_col, int k, int e, const TX *xb)
{
    asm volatile("" : "+l"(xb));
    TX s = 0;
    while (k + (B - 1) * G < e) {
        TX xv[B];
#pragma unroll
        for (int u = 0; u < B; ++u) xv[u] = ld_xg<COH>(xb + s_col[k + u * G]);
#pragma unroll
        for (int u = 0; u < B; ++u) s += (TX)s_val[k + u * G] * xv[u];
        k += B * G;
    }
    if (k < e) {
        TX xv[B];
#pragma unroll
        for (int u = 0; u < B; ++u) xv[u] = (k + u * G < e) ? ld_xg<COH>(xb + s_col[k + u * G]) : (TX)0;
#pragma unroll
        for (int u = 0; u < B; ++u)
            if (k + u * G < e) s += (TX)s_val[k + u * G] * xv[u];
    }
    return s;
}

// LIMIT > 0: rows longer than LIMIT entries are left to reduce_rows_long (a thread walking a 64-entry row of a power-law
// matrix alone needs eight dependent rounds of gathers and is what its CTA -- and the tail of the product -- waits for)
template <typename TV, typename TX, int THREADS, int B, bool COH = false, int LIMIT = 0>
__device__ __forceinline__ void reduce_rows_lean2(const TV *s_val, const int *s_col, const int *s_rp, int sb, int nrows,
                                                  int row_glob, const TX *__restrict__ x, int base, TX *__restrict__ y,
                                                  TX alpha, TX beta)
{
    const TX *xb = x - base;
    if (nrows > THREADS) {
        // more rows than threads (short rows: config 4): a thread's two rows go TOGETHER when both fit one batch -- their gathers
        // are issued back to back, one round trip instead of two (each row is still summed left to right)
        asm volatile("" : "+l"(xb));
        for (int i = threadIdx.x; i < nrows; i += 2 * THREADS) {
            const int i2 = i + THREADS;
            const int k0 = s_rp[i] - sb, e0 = s_rp[i + 1] - sb;
            bool two = i2 < nrows;
            int k1 = two ? s_rp[i2] - sb : 0, e1 = two ? s_rp[i2 + 1] - sb : 0;
            if (LIMIT > 0) {
                if (two && e1 - k1 > LIMIT) { two = false; k1 = e1 = 0; }
                if (e0 - k0 > LIMIT) { // only the second row (if any) is this thread's
                    if (two) store_y(y, row_glob + i2, row_dot_lean2<TV, TX, 1, B, COH>(s_val, s_col, k1, e1, xb), alpha, beta);
                    continue;
                }
            }
            if (e0 - k0 <= B && e1 - k1 <= B) {
                TX xv0[B], xv1[B];
#pragma unroll
                for (int u = 0; u < B; ++u) xv0[u] = (k0 + u < e0) ? ld_xg<COH>(xb + s_col[k0 + u]) : (TX)0;
#pragma unroll
                for (int u = 0; u < B; ++u) xv1[u] = (k1 + u < e1) ? ld_xg<COH>(xb + s_col[k1 + u]) : (TX)0;
                TX s0 = 0, s1 = 0;
#pragma unroll
                for (int u = 0; u < B; ++u)
                    if (k0 + u < e0) s0 += (TX)s_val[k0 + u] * xv0[u];
#pragma unroll
                for (int u = 0; u < B; ++u)
                    if (k1 + u < e1) s1 += (TX)s_val[k1 + u] * xv1[u];
                store_y(y, row_glob + i, s0, alpha, beta);
                if (two) store_y(y, row_glob + i2, s1, alpha, beta);
            } else {
                store_y(y, row_glob + i, row_dot_lean2<TV, TX, 1, B, COH>(s_val, s_col, k0, e0, xb), alpha, beta);
                if (two) store_y(y, row_glob + i2, row_dot_lean2<TV, TX, 1, B, COH>(s_val, s_col, k1, e1, xb), alpha, beta);
            }
        }
        return;
    }
    for (int i = threadIdx.x; i < nrows; i += THREADS) {
        if (LIMIT > 0 && s_rp[i + 1] - s_rp[i] > LIMIT) continue;
        const TX s = row_dot_lean2<TV, TX, 1, B, COH>(s_val, s_col, s_rp[i] - sb, s_rp[i + 1] - sb, xb);
        store_y(y, row_glob + i, s, alpha, beta);
    }
}

// The rows reduce_rows_lean2<..., LIMIT> left out: G lanes each (every group scans its share of the tile's rows)
template <typename TV, typename TX, int THREADS, int G, int B, bool COH, int LIMIT>
__device__ __forceinline__ void reduce_rows_long(const TV *s_val, const int *s_col, const int *s_rp, int sb, int nrows,
                                                 int row_glob, const TX *__restrict__ x, int base, TX *__restrict__ y,
                                                 TX alpha, TX beta)
{
    constexpr int GROUPS = THREADS / G;
    const int gid = threadIdx.x / G, lane = threadIdx.x % G;
    const TX *xb = x - base;
    const int iters = (nrows + GROUPS - 1) / GROUPS;
    for (int it = 0; it < iters; ++it) {
        const int i = it * GROUPS + gid;
        const bool lg = i < nrows && s_rp[i + 1] - s_rp[i] > LIMIT;
        if (!__any_sync(0xffffffffu, lg)) continue;
        TX s = 0;
        if (lg) s = row_dot_lean2<TV, TX, G, B, COH>(s_val, s_col, s_rp[i] - sb + lane, s_rp[i + 1] - sb, xb);
#pragma unroll
        for (int o = G / 2; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        if (lg && lane == 0) store_y(y, row_glob + i, s, alpha, beta);
    }
}

// Product staging for scattered matrices: the CTA turns the staged values into products a_k * x[col_k] IN PLACE, every thread
// taking entries tid, tid + THREADS, ... with eight gathers in flight -- ONE round of gathers for the whole tile, whatever its row
// lengths (thread-per-row needs a dependent round per eight entries of its longest row: 3.5 us median per tile on config 4,
// of which the arithmetic is nothing).  The row sums are then shared-memory additions in stored order: the same bits as
// multiply-then-add in a thread (the library is compiled without contraction).  Needs sizeof(TV) == sizeof(TX).
template <typename TV, typename TX, int THREADS, bool COH>
__device__ __forceinline__ void stage_products(TV *s_val, const int *s_col, int kb, int ke, const TX *__restrict__ x, int base)
{
    static_assert(sizeof(TV) == sizeof(TX), "products are staged in place");
    const TX *xb = x - base;
    asm volatile("" : "+l"(xb));
    int k = kb + (int)threadIdx.x;
    for (; k + 7 * THREADS < ke; k += 8 * THREADS) {
        TX xv[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) xv[u] = ld_xg<COH>(xb + s_col[k + u * THREADS]);
#pragma unroll
        for (int u = 0; u < 8; ++u) s_val[k + u * THREADS] = (TV)((TX)s_val[k + u * THREADS] * xv[u]);
    }
    {
        TX xv[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) xv[u] = (k + u * THREADS < ke) ? ld_xg<COH>(xb + s_col[k + u * THREADS]) : (TX)0;
#pragma unroll
        for (int u = 0; u < 8; ++u)
            if (k + u * THREADS < ke) s_val[k + u * THREADS] = (TV)((TX)s_val[k + u * THREADS] * xv[u]);
    }
}

// rows of staged products: thread per row up to LIMIT entries (stored order), G lanes per longer row
template <typename TV, typename TX, int THREADS, int G, int LIMIT>
__device__ __forceinline__ void sum_staged_rows(const TV *s_val, const int *s_rp, int sb, int nrows, int row_glob, TX *__restrict__ y,
                                                TX alpha, TX beta, bool any_long)
{
    for (int i = threadIdx.x; i < nrows; i += THREADS) {
        const int k0 = s_rp[i] - sb, e = s_rp[i + 1] - sb;
        if (e - k0 > LIMIT) continue;
        TX s = 0;
        for (int k = k0; k < e; ++k) s += (TX)s_val[k];
        store_y(y, row_glob + i, s, alpha, beta);
    }
    if (!any_long) return;
    constexpr int GROUPS = THREADS / G;
    const int gid = threadIdx.x / G, lane = threadIdx.x % G;
    const int iters = (nrows + GROUPS - 1) / GROUPS;
    for (int it = 0; it < iters; ++it) {
        const int i = it * GROUPS + gid;
        const bool lg = i < nrows && s_rp[i + 1] - s_rp[i] > LIMIT;
        if (!__any_sync(0xffffffffu, lg)) continue;
        TX s = 0;
        if (lg)
            for (int k = s_rp[i] - sb + lane, e = s_rp[i + 1] - sb; k < e; k += G) s += (TX)s_val[k];
#pragma unroll
        for (int o = G / 2; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        if (lg && lane == 0) store_y(y, row_glob + i, s, alpha, beta);
    }
}

// G lanes per row with the same unpredicated batches (lane l takes entries l, l+G, ...)
template <typename TV, typename TX, int THREADS, int G, int B, bool COH = false>
__device__ __forceinline__ void reduce_rows_lean2g(const TV *s_val, const int *s_col, const int *s_rp, int sb, int nrows,
                                                   int row_glob, const TX *__restrict__ x, int base, TX *__restrict__ y,
                                                   TX alpha, TX beta)
{
    constexpr int GROUPS = THREADS / G;
    const int gid = threadIdx.x / G, lane = threadIdx.x % G;
    const TX *xb = x - base;
    const int iters = (nrows + GROUPS - 1) / GROUPS;
    for (int it = 0; it < iters; ++it) {
        const int i = it * GROUPS + gid;
        TX s = 0;
        if (i < nrows) s = row_dot_lean2<TV, TX, G, B, COH>(s_val, s_col, s_rp[i] - sb + lane, s_rp[i + 1] - sb, xb);
#pragma unroll
        for (int o = G / 2; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        if (i < nrows && lane == 0) store_y(y, row_glob + i, s, alpha, beta);
    }
}

template <typename TV, typename TX, int THREADS, int G, bool COH = false>
__device__ __forceinline__ void reduce_rows_lean(const TV *s_val, const int *s_col, const int *s_rp, int sb, int nrows,
                                                 int row_glob, const TX *__restrict__ x, int base, TX *__restrict__ y,
                                                 TX alpha, TX beta)
{
    const int tid = threadIdx.x;
    const TX *xb = x - base;
    if (G == 1) {
        for (int i = tid; i < nrows; i += THREADS) {
            const TX s = row_dot_lean<TV, TX, 1, 8, COH>(s_val, s_col, s_rp[i] - sb, s_rp[i + 1] - sb, xb);
            store_y(y, row_glob + i, s, alpha, beta);
        }
    } else {
        constexpr int GROUPS = THREADS / G;
        const int gid = tid / G, lane = tid % G;
        const int iters = (nrows + GROUPS - 1) / GROUPS;
        for (int it = 0; it < iters; ++it) {
            const int i = it * GROUPS + gid;
            TX s = 0;
            if (i < nrows) s = row_dot_lean<TV, TX, G, 4, COH>(s_val, s_col, s_rp[i] - sb + lane, s_rp[i + 1] - sb, xb);
#pragma unroll
            for (int o = G / 2; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
            if (i < nrows && lane == 0) store_y(y, row_glob + i, s, alpha, beta);
        }
    }
}

} // namespace rsg
