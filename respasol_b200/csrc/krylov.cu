// krylov.cu -- preconditioned conjugate gradients (SPD) and BiCGStab (general) on the three hot-path calls
// (SURVEY.md 8f-2): the loop the
// reference benchmarks its kernels for (mixed-precision sparse solvers, /root/reference/README.md:5-7) but does
// not contain.  q = A p is spmv_csr_d; z = M^-1 r is sptrsv_lu_d on fp64 ILU(0) factors, or sptrsv_lu_s on fp32
// factors with r rounded to fp32 and z widened back (fp32 preconditioner inside an fp64 iteration); the vector
// updates and dot products are fused into three small kernels whose scalars (alpha, beta, r.z, p.q) never leave
// the device.  Dot products are deterministic: per-CTA partials, the last CTA to arrive adds them in CTA order.
//
// One iteration = 1 SpMV + 2 triangular solves + 3 (4 with fp32 factors) vector kernels; the host reads one
// double (|r|^2) per iteration to decide convergence.
#include <cuda_runtime.h>

#include <cmath>

#include "rsg_internal.cuh"

namespace rsg {

static constexpr int kVecThreads = 256;
static constexpr int kMaxVecBlocks = 1024;

// scalars (device doubles): [0] rz (ping), [1] rz (pong), [2] pq, [3] rr, [4] bb
enum { S_RZ0 = 0, S_RZ1 = 1, S_PQ = 2, S_RR = 3, S_BB = 4, S_COUNT = 8 };

// CTA partial -> partials[blockIdx]; the last CTA adds all partials in index order and stores *out.
__device__ __forceinline__ void finish_dot(double local, double *partials, unsigned *ticket, double *out)
{
    __shared__ double s_red[kVecThreads / 32];
    __shared__ bool s_last;
    const int tid = threadIdx.x;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) local += __shfl_xor_sync(0xffffffffu, local, o);
    if ((tid & 31) == 0) s_red[tid >> 5] = local;
    __syncthreads();
    if (tid == 0) {
        double t = 0;
        for (int w = 0; w < kVecThreads / 32; ++w) t += s_red[w];
        partials[blockIdx.x] = t;
        __threadfence();
        s_last = atomicAdd(ticket, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    double t = 0;
    for (int i = tid; i < (int)gridDim.x; i += kVecThreads) t += *reinterpret_cast<volatile double *>(partials + i);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
    __syncthreads();
    if ((tid & 31) == 0) s_red[tid >> 5] = t;
    __syncthreads();
    if (tid == 0) {
        double s = 0;
        for (int w = 0; w < kVecThreads / 32; ++w) s += s_red[w];
        *out = s;
        *ticket = 0;
    }
}

// out = a . b
__global__ void __launch_bounds__(kVecThreads) dot_kernel(int n, const double *__restrict__ a, const double *__restrict__ b,
                                                          double *partials, unsigned *ticket, double *out)
{
    double s = 0;
    for (int i = blockIdx.x * kVecThreads + threadIdx.x; i < n; i += gridDim.x * kVecThreads) s = fma(a[i], b[i], s);
    finish_dot(s, partials, ticket, out);
}

// r = b - q (q = A x0), bb = b.b, rr = r.r in one pass (two dots: bb first into partials, rr second)
__global__ void __launch_bounds__(kVecThreads) residual_kernel(int n, const double *__restrict__ b, const double *__restrict__ q,
                                                               double *__restrict__ r)
{
    for (int i = blockIdx.x * kVecThreads + threadIdx.x; i < n; i += gridDim.x * kVecThreads) r[i] = b[i] - q[i];
}

// alpha = rz / pq;  x += alpha p;  r -= alpha q;  rr = r.r;  (r32 = (float) r when the preconditioner is fp32)
__global__ void __launch_bounds__(kVecThreads) pcg_update_xr_kernel(int n, const double *scal, int rz_slot, const double *__restrict__ p,
                                                                    const double *__restrict__ q, double *__restrict__ x,
                                                                    double *__restrict__ r, float *__restrict__ r32, double *partials,
                                                                    unsigned *ticket, double *out_rr)
{
    const double alpha = scal[rz_slot] / scal[S_PQ];
    double s = 0;
    for (int i = blockIdx.x * kVecThreads + threadIdx.x; i < n; i += gridDim.x * kVecThreads) {
        x[i] = fma(alpha, p[i], x[i]);
        const double ri = fma(-alpha, q[i], r[i]);
        r[i] = ri;
        if (r32) r32[i] = (float)ri;
        s = fma(ri, ri, s);
    }
    finish_dot(s, partials, ticket, out_rr);
}

// z = (double) z32 (fp32 preconditioner), rz_new = r.z
__global__ void __launch_bounds__(kVecThreads) pcg_rz_kernel(int n, const double *__restrict__ r, const float *__restrict__ z32,
                                                             double *__restrict__ z, double *partials, unsigned *ticket, double *out)
{
    double s = 0;
    for (int i = blockIdx.x * kVecThreads + threadIdx.x; i < n; i += gridDim.x * kVecThreads) {
        double zi;
        if (z32) { zi = (double)z32[i]; z[i] = zi; } else zi = z[i];
        s = fma(r[i], zi, s);
    }
    finish_dot(s, partials, ticket, out);
}

// beta = rz_new / rz_old (beta = 0 on the first pass: p = z);  p = z + beta p
__global__ void __launch_bounds__(kVecThreads) pcg_update_p_kernel(int n, const double *scal, int rz_new_slot, int first,
                                                                   const double *__restrict__ z, double *__restrict__ p)
{
    const double beta = first ? 0.0 : scal[rz_new_slot] / scal[rz_new_slot ^ 1];
    for (int i = blockIdx.x * kVecThreads + threadIdx.x; i < n; i += gridDim.x * kVecThreads) p[i] = first ? z[i] : fma(beta, p[i], z[i]);
}

__global__ void __launch_bounds__(kVecThreads) cast_kernel(int n, const double *__restrict__ a, float *__restrict__ b)
{
    for (int i = blockIdx.x * kVecThreads + threadIdx.x; i < n; i += gridDim.x * kVecThreads) b[i] = (float)a[i];
}

// ---- BiCGStab (van der Vorst 1992, right-preconditioned).  scalars: [0],[1] rho ping-pong, [2] rh.v, [3] rr, [4] bb,
// [5] t.s, [6] t.t.  alpha = rho/rh.v and omega = t.s/t.t are recomputed from them by every kernel that needs them.
enum { B_RHV = 2, B_RR = 3, B_BB = 4, B_TS = 5, B_TT = 6 };

// p = r + beta (p - omega v), beta = (rho_new/rho_old)(alpha/omega); first pass: p = r.  (p32 = (float) p)
__global__ void __launch_bounds__(kVecThreads) bicg_p_kernel(int n, const double *scal, int rho_slot, int first, const double *__restrict__ r,
                                                             const double *__restrict__ v, double *__restrict__ p, float *__restrict__ p32)
{
    double beta = 0, omega = 0;
    if (!first) {
        const double rho_new = scal[rho_slot], rho_old = scal[rho_slot ^ 1];
        const double alpha = rho_old / scal[B_RHV];
        omega = scal[B_TS] / scal[B_TT];
        beta = (rho_new / rho_old) * (alpha / omega);
    }
    for (int i = blockIdx.x * kVecThreads + threadIdx.x; i < n; i += gridDim.x * kVecThreads) {
        const double pi = first ? r[i] : fma(beta, fma(-omega, v[i], p[i]), r[i]);
        p[i] = pi;
        if (p32) p32[i] = (float)pi;
    }
}

// s = r - alpha v (in place on r), alpha = rho/rh.v;  (s32 = (float) s)
__global__ void __launch_bounds__(kVecThreads) bicg_s_kernel(int n, const double *scal, int rho_slot, const double *__restrict__ v,
                                                             double *__restrict__ r, float *__restrict__ s32)
{
    const double alpha = scal[rho_slot] / scal[B_RHV];
    for (int i = blockIdx.x * kVecThreads + threadIdx.x; i < n; i += gridDim.x * kVecThreads) {
        const double si = fma(-alpha, v[i], r[i]);
        r[i] = si;
        if (s32) s32[i] = (float)si;
    }
}

// out0 = a.b, out1 = a.a
__global__ void __launch_bounds__(kVecThreads) dot2_kernel(int n, const double *__restrict__ a, const double *__restrict__ b,
                                                           double *partials, unsigned *ticket, double *out0, double *out1)
{
    double s0 = 0, s1 = 0;
    for (int i = blockIdx.x * kVecThreads + threadIdx.x; i < n; i += gridDim.x * kVecThreads) {
        const double ai = a[i];
        s0 = fma(ai, b[i], s0);
        s1 = fma(ai, ai, s1);
    }
    finish_dot(s0, partials, ticket, out0);
    __syncthreads();
    finish_dot(s1, partials + kMaxVecBlocks, ticket + 1, out1);
}

// x += alpha y + omega z;  r = s - omega t (s lives in r);  rr = r.r;  rho_next = rh.r
__global__ void __launch_bounds__(kVecThreads) bicg_xr_kernel(int n, const double *scal, int rho_slot, const double *__restrict__ y,
                                                              const double *__restrict__ z, const double *__restrict__ t,
                                                              const double *__restrict__ rh, double *__restrict__ x, double *__restrict__ r,
                                                              double *partials, unsigned *ticket, double *out_rr, double *out_rho)
{
    const double alpha = scal[rho_slot] / scal[B_RHV], omega = scal[B_TS] / scal[B_TT];
    double s0 = 0, s1 = 0;
    for (int i = blockIdx.x * kVecThreads + threadIdx.x; i < n; i += gridDim.x * kVecThreads) {
        x[i] = fma(omega, z[i], fma(alpha, y[i], x[i]));
        const double ri = fma(-omega, t[i], r[i]);
        r[i] = ri;
        s0 = fma(ri, ri, s0);
        s1 = fma(rh[i], ri, s1);
    }
    finish_dot(s0, partials, ticket, out_rr);
    __syncthreads();
    finish_dot(s1, partials + kMaxVecBlocks, ticket + 1, out_rho);
}

__global__ void __launch_bounds__(kVecThreads) widen_kernel(int n, const float *__restrict__ a, double *__restrict__ b)
{
    for (int i = blockIdx.x * kVecThreads + threadIdx.x; i < n; i += gridDim.x * kVecThreads) b[i] = (double)a[i];
}

// ---- GMRES(30): modified Gram-Schmidt step.  w -= h_prev v_prev (when v_prev);  out = w . v_next (v_next == w: |w|^2).
// One launch per basis vector: the subtraction of vector i-1 rides on the pass that takes the dot with vector i.
// (w and v_next may alias: no __restrict__ on them.)
__global__ void __launch_bounds__(kVecThreads) gmres_mgs_kernel(int n, const double *h_prev, const double *v_prev, const double *v_next,
                                                                double *w, double *partials, unsigned *ticket, double *out)
{
    const double hp = v_prev ? *h_prev : 0.0;
    const bool norm = v_next == w;
    double s = 0;
    for (int i = blockIdx.x * kVecThreads + threadIdx.x; i < n; i += gridDim.x * kVecThreads) {
        double wi = w[i];
        if (v_prev) { wi = fma(-hp, v_prev[i], wi); w[i] = wi; }
        s = fma(wi, norm ? wi : v_next[i], s);
    }
    finish_dot(s, partials, ticket, out);
}

// v = w / sqrt(*ww)  (v32 = (float) v; v may be w)
__global__ void __launch_bounds__(kVecThreads) gmres_scale_kernel(int n, const double *ww, const double *w, double *v, float *__restrict__ v32)
{
    const double inv = 1.0 / sqrt(*ww);
    for (int i = blockIdx.x * kVecThreads + threadIdx.x; i < n; i += gridDim.x * kVecThreads) {
        const double vi = w[i] * inv;
        v[i] = vi;
        if (v32) v32[i] = (float)vi;
    }
}

// x += sum_j y[j] Z_j
__global__ void __launch_bounds__(kVecThreads) gmres_update_x_kernel(int n, int k, const double *__restrict__ y, const double *__restrict__ Z,
                                                                     size_t ld, double *__restrict__ x)
{
    for (int i = blockIdx.x * kVecThreads + threadIdx.x; i < n; i += gridDim.x * kVecThreads) {
        double s = 0;
        for (int j = 0; j < k; ++j) s = fma(y[j], Z[(size_t)j * ld + i], s);
        x[i] += s;
    }
}

} // namespace rsg

using namespace rsg;

namespace {
constexpr size_t kWsTail = 2 * kMaxVecBlocks * 8 + S_COUNT * 8 + 64; // partial sums (two dots), scalars, tickets

int krylov_workspace(rsg_context *h, size_t bytes)
{
    if (h->krylov_bytes < bytes) {
        if (h->krylov_ws) cudaFree(h->krylov_ws);
        h->krylov_ws = nullptr;
        h->krylov_bytes = 0;
        if (cudaMalloc(&h->krylov_ws, bytes) != cudaSuccess) { (void)cudaGetLastError(); return RSG_ALLOC_FAILED; }
        h->krylov_bytes = bytes;
    }
    if (!h->krylov_pin && cudaMallocHost(&h->krylov_pin, (S_COUNT + 2 * 30 + 8) * sizeof(double)) != cudaSuccess) { (void)cudaGetLastError(); return RSG_ALLOC_FAILED; }
    return RSG_OK;
}

struct Guard {
    int prev = 0;
    explicit Guard(int dev) { cudaGetDevice(&prev); if (prev != dev) cudaSetDevice(dev); else prev = -1; }
    ~Guard() { if (prev >= 0) cudaSetDevice(prev); }
};
} // namespace

extern "C" int rsg_pcg_d(rsg_handle_t h, int n, int nnz, int base, const int *rowptr, const int *colidx, const double *values,
                         const void *lu_values, int lu_precision, const double *b, double *x, int max_iters, double rtol,
                         int *iters, double *relres)
{
    if (!h || n < 0 || !rowptr || (n > 0 && (!b || !x || !values)) || max_iters < 0 ||
        (lu_precision != 0 && lu_precision != 'd' && lu_precision != 's') || (lu_precision != 0 && !lu_values))
        return RSG_BAD_ARG;
    if (iters) *iters = 0;
    if (relres) *relres = 0.0;
    if (n == 0) return RSG_OK;
    Guard g(h->device);
    cudaStream_t st = h->stream;
    const bool pre = lu_precision != 0, pre32 = lu_precision == 's';
    // workspace: r, z, p, q (fp64), r32, z32, t32 (fp32), partials, scalars, ticket
    const size_t nd = ((size_t)n + 3) & ~(size_t)3;
    RSG_TRY(krylov_workspace(h, nd * 8 * 4 + nd * 4 * 3 + kWsTail));
    char *w = static_cast<char *>(h->krylov_ws);
    double *r = reinterpret_cast<double *>(w); w += nd * 8;
    double *z = reinterpret_cast<double *>(w); w += nd * 8;
    double *p = reinterpret_cast<double *>(w); w += nd * 8;
    double *q = reinterpret_cast<double *>(w); w += nd * 8;
    float *r32 = reinterpret_cast<float *>(w); w += nd * 4;
    float *z32 = reinterpret_cast<float *>(w); w += nd * 4;
    float *t32 = reinterpret_cast<float *>(w); w += nd * 4;
    double *partials = reinterpret_cast<double *>(w); w += 2 * kMaxVecBlocks * 8;
    double *scal = reinterpret_cast<double *>(w); w += S_COUNT * 8;
    unsigned *ticket = reinterpret_cast<unsigned *>(w);
    RSG_CUDA_TRY(cudaMemsetAsync(scal, 0, S_COUNT * 8 + 64, st));
    const int grid = std::max(1, std::min(kMaxVecBlocks, std::min(h->sm_count * 8, (n + kVecThreads * 4 - 1) / (kVecThreads * 4))));
    double *pin = h->krylov_pin;

    auto precondition = [&](bool first) -> int { // z = M^-1 r, rz -> scal[slot]
        (void)first;
        if (!pre) { RSG_CUDA_TRY(cudaMemcpyAsync(z, r, sizeof(double) * (size_t)n, cudaMemcpyDeviceToDevice, st)); return RSG_OK; }
        if (pre32) return sptrsv_lu_s(h, n, nnz, base, rowptr, colidx, static_cast<const float *>(lu_values), r32, t32, z32);
        return sptrsv_lu_d(h, n, nnz, base, rowptr, colidx, static_cast<const double *>(lu_values), r, q, z); // q is free here
    };

    // r = b - A x0;  bb = b.b;  rr = r.r
    RSG_TRY(spmv_csr_d(h, n, n, nnz, base, rowptr, colidx, values, 1.0, x, 0.0, q));
    residual_kernel<<<grid, kVecThreads, 0, st>>>(n, b, q, r);
    dot_kernel<<<grid, kVecThreads, 0, st>>>(n, b, b, partials, ticket, scal + S_BB);
    dot_kernel<<<grid, kVecThreads, 0, st>>>(n, r, r, partials, ticket, scal + S_RR);
    h->launches += 3;
    if (pre32) { cast_kernel<<<grid, kVecThreads, 0, st>>>(n, r, r32); h->launches++; }
    RSG_CUDA_TRY(cudaMemcpyAsync(pin, scal, S_COUNT * sizeof(double), cudaMemcpyDeviceToHost, st));
    RSG_CUDA_TRY(cudaStreamSynchronize(st));
    const double bb = pin[S_BB] > 0 ? pin[S_BB] : 1.0;
    double rr = pin[S_RR];
    int it = 0, slot = S_RZ0;
    const double stop = rtol * rtol * bb;
    if (!(rr <= stop)) {
        RSG_TRY(precondition(true));
        pcg_rz_kernel<<<grid, kVecThreads, 0, st>>>(n, r, pre32 ? z32 : nullptr, z, partials, ticket, scal + slot);
        pcg_update_p_kernel<<<grid, kVecThreads, 0, st>>>(n, scal, slot, 1, z, p);
        h->launches += 2;
        while (it < max_iters) {
            RSG_TRY(spmv_csr_d(h, n, n, nnz, base, rowptr, colidx, values, 1.0, p, 0.0, q));
            dot_kernel<<<grid, kVecThreads, 0, st>>>(n, p, q, partials, ticket, scal + S_PQ);
            pcg_update_xr_kernel<<<grid, kVecThreads, 0, st>>>(n, scal, slot, p, q, x, r, pre32 ? r32 : nullptr, partials, ticket, scal + S_RR);
            h->launches += 2;
            RSG_CUDA_TRY(cudaMemcpyAsync(pin + S_RR, scal + S_RR, sizeof(double), cudaMemcpyDeviceToHost, st));
            ++it;
            // the next preconditioner application is issued before the host looks at |r|: the stream stays busy
            RSG_TRY(precondition(false));
            slot ^= 1;
            pcg_rz_kernel<<<grid, kVecThreads, 0, st>>>(n, r, pre32 ? z32 : nullptr, z, partials, ticket, scal + slot);
            pcg_update_p_kernel<<<grid, kVecThreads, 0, st>>>(n, scal, slot, 0, z, p);
            h->launches += 2;
            RSG_CUDA_TRY(cudaStreamSynchronize(st));
            rr = pin[S_RR];
            if (!(rr > stop) || !std::isfinite(rr)) break;
        }
    }
    RSG_CUDA_TRY(cudaGetLastError());
    if (iters) *iters = it;
    if (relres) *relres = std::sqrt(rr / bb);
    if (!std::isfinite(rr)) return check_err_flag(h) == RSG_DEPENDENCY_TIMEOUT ? RSG_DEPENDENCY_TIMEOUT : RSG_BREAKDOWN; // a hung solve leaves sentinel NaNs
    return RSG_OK;
}

extern "C" int rsg_bicgstab_d(rsg_handle_t h, int n, int nnz, int base, const int *rowptr, const int *colidx, const double *values,
                              const void *lu_values, int lu_precision, const double *b, double *x, int max_iters, double rtol,
                              int *iters, double *relres)
{
    if (!h || n < 0 || !rowptr || (n > 0 && (!b || !x || !values)) || max_iters < 0 ||
        (lu_precision != 0 && lu_precision != 'd' && lu_precision != 's') || (lu_precision != 0 && !lu_values))
        return RSG_BAD_ARG;
    if (iters) *iters = 0;
    if (relres) *relres = 0.0;
    if (n == 0) return RSG_OK;
    Guard g(h->device);
    cudaStream_t st = h->stream;
    const bool pre = lu_precision != 0, pre32 = lu_precision == 's';
    const size_t nd = ((size_t)n + 3) & ~(size_t)3;
    RSG_TRY(krylov_workspace(h, nd * 8 * 7 + nd * 4 * 3 + kWsTail));
    char *w = static_cast<char *>(h->krylov_ws);
    double *r = reinterpret_cast<double *>(w); w += nd * 8;  // r, then s in place
    double *rh = reinterpret_cast<double *>(w); w += nd * 8; // shadow residual
    double *p = reinterpret_cast<double *>(w); w += nd * 8;
    double *v = reinterpret_cast<double *>(w); w += nd * 8;
    double *y = reinterpret_cast<double *>(w); w += nd * 8;  // K^-1 p
    double *z = reinterpret_cast<double *>(w); w += nd * 8;  // K^-1 s
    double *t = reinterpret_cast<double *>(w); w += nd * 8;  // A z; scratch of the fp64 triangular solves before that
    float *in32 = reinterpret_cast<float *>(w); w += nd * 4;
    float *out32 = reinterpret_cast<float *>(w); w += nd * 4;
    float *tmp32 = reinterpret_cast<float *>(w); w += nd * 4;
    double *partials = reinterpret_cast<double *>(w); w += 2 * kMaxVecBlocks * 8;
    double *scal = reinterpret_cast<double *>(w); w += S_COUNT * 8;
    unsigned *ticket = reinterpret_cast<unsigned *>(w);
    RSG_CUDA_TRY(cudaMemsetAsync(scal, 0, S_COUNT * 8 + 64, st));
    const int grid = std::max(1, std::min(kMaxVecBlocks, std::min(h->sm_count * 8, (n + kVecThreads * 4 - 1) / (kVecThreads * 4))));
    double *pin = h->krylov_pin;

    // out = K^-1 in (in32 already holds (float) in when the factors are fp32)
    auto precondition = [&](const double *in, double *out) -> int {
        if (!pre) { RSG_CUDA_TRY(cudaMemcpyAsync(out, in, sizeof(double) * (size_t)n, cudaMemcpyDeviceToDevice, st)); return RSG_OK; }
        if (pre32) {
            RSG_TRY(sptrsv_lu_s(h, n, nnz, base, rowptr, colidx, static_cast<const float *>(lu_values), in32, tmp32, out32));
            widen_kernel<<<grid, kVecThreads, 0, st>>>(n, out32, out);
            h->launches++;
            return RSG_OK;
        }
        return sptrsv_lu_d(h, n, nnz, base, rowptr, colidx, static_cast<const double *>(lu_values), in, t, out);
    };

    RSG_TRY(spmv_csr_d(h, n, n, nnz, base, rowptr, colidx, values, 1.0, x, 0.0, v));
    residual_kernel<<<grid, kVecThreads, 0, st>>>(n, b, v, r);
    RSG_CUDA_TRY(cudaMemcpyAsync(rh, r, sizeof(double) * (size_t)n, cudaMemcpyDeviceToDevice, st));
    dot_kernel<<<grid, kVecThreads, 0, st>>>(n, b, b, partials, ticket, scal + B_BB);
    int slot = 0;
    dot2_kernel<<<grid, kVecThreads, 0, st>>>(n, r, r, partials, ticket, scal + slot, scal + B_RR); // rho = rh.r = r.r
    h->launches += 3;
    RSG_CUDA_TRY(cudaMemcpyAsync(pin, scal, S_COUNT * sizeof(double), cudaMemcpyDeviceToHost, st));
    RSG_CUDA_TRY(cudaStreamSynchronize(st));
    const double bb = pin[B_BB] > 0 ? pin[B_BB] : 1.0;
    double rr = pin[B_RR];
    const double stop = rtol * rtol * bb;
    int it = 0;
    while (it < max_iters && rr > stop && std::isfinite(rr)) {
        bicg_p_kernel<<<grid, kVecThreads, 0, st>>>(n, scal, slot, it == 0, r, v, p, pre32 ? in32 : nullptr);
        RSG_TRY(precondition(p, y));
        RSG_TRY(spmv_csr_d(h, n, n, nnz, base, rowptr, colidx, values, 1.0, y, 0.0, v));
        dot_kernel<<<grid, kVecThreads, 0, st>>>(n, rh, v, partials, ticket, scal + B_RHV);
        bicg_s_kernel<<<grid, kVecThreads, 0, st>>>(n, scal, slot, v, r, pre32 ? in32 : nullptr);
        RSG_TRY(precondition(r, z));
        RSG_TRY(spmv_csr_d(h, n, n, nnz, base, rowptr, colidx, values, 1.0, z, 0.0, t));
        dot2_kernel<<<grid, kVecThreads, 0, st>>>(n, t, r, partials, ticket, scal + B_TS, scal + B_TT);
        bicg_xr_kernel<<<grid, kVecThreads, 0, st>>>(n, scal, slot, y, z, t, rh, x, r, partials, ticket, scal + B_RR, scal + (slot ^ 1));
        h->launches += 5;
        slot ^= 1;
        RSG_CUDA_TRY(cudaMemcpyAsync(pin + B_RR, scal + B_RR, sizeof(double), cudaMemcpyDeviceToHost, st));
        RSG_CUDA_TRY(cudaStreamSynchronize(st));
        rr = pin[B_RR];
        ++it;
    }
    RSG_CUDA_TRY(cudaGetLastError());
    if (iters) *iters = it;
    if (relres) *relres = std::sqrt(rr / bb);
    if (!std::isfinite(rr)) return check_err_flag(h) == RSG_DEPENDENCY_TIMEOUT ? RSG_DEPENDENCY_TIMEOUT : RSG_BREAKDOWN; // a hung solve leaves sentinel NaNs
    return RSG_OK;
}

// Restarted flexible GMRES(kGmresRestart), right-preconditioned (Saad & Schultz 1986; Saad Alg. 9.6): the preconditioned
// vectors z_j = K^-1 v_j are kept and x += Z y, so fp32 factors (whose input is rounded) cost no accuracy.  Both bases stay in HBM; the Hessenberg
// column of one Arnoldi step is built on the device by j+2 fused subtract-and-dot launches and read back in ONE copy; the
// Givens rotations and the small triangular solve run on the host.
static constexpr int kGmresRestart = 30;

extern "C" int rsg_gmres_d(rsg_handle_t h, int n, int nnz, int base, const int *rowptr, const int *colidx, const double *values,
                           const void *lu_values, int lu_precision, const double *b, double *x, int max_iters, double rtol,
                           int *iters, double *relres)
{
    if (!h || n < 0 || !rowptr || (n > 0 && (!b || !x || !values)) || max_iters < 0 ||
        (lu_precision != 0 && lu_precision != 'd' && lu_precision != 's') || (lu_precision != 0 && !lu_values))
        return RSG_BAD_ARG;
    if (iters) *iters = 0;
    if (relres) *relres = 0.0;
    if (n == 0) return RSG_OK;
    constexpr int R = kGmresRestart;
    Guard g(h->device);
    cudaStream_t st = h->stream;
    const bool pre = lu_precision != 0, pre32 = lu_precision == 's';
    const size_t nd = ((size_t)n + 3) & ~(size_t)3;
    // workspace: bases V (R+1 vectors) and Z (R), w, t (fp64), three fp32 vectors, Hessenberg column, y, partials, scalars, ticket
    RSG_TRY(krylov_workspace(h, nd * 8 * (2 * R + 3) + nd * 4 * 3 + (size_t)(2 * R + 4) * 8 + kWsTail));
    char *w_ = static_cast<char *>(h->krylov_ws);
    double *V = reinterpret_cast<double *>(w_); w_ += nd * 8 * (R + 1);
    double *w = reinterpret_cast<double *>(w_); w_ += nd * 8;
    double *Z = reinterpret_cast<double *>(w_); w_ += nd * 8 * R;
    double *t = reinterpret_cast<double *>(w_); w_ += nd * 8;
    float *in32 = reinterpret_cast<float *>(w_); w_ += nd * 4;
    float *out32 = reinterpret_cast<float *>(w_); w_ += nd * 4;
    float *tmp32 = reinterpret_cast<float *>(w_); w_ += nd * 4;
    double *hcol = reinterpret_cast<double *>(w_); w_ += (R + 2) * 8;
    double *ycoef = reinterpret_cast<double *>(w_); w_ += (R + 2) * 8;
    double *partials = reinterpret_cast<double *>(w_); w_ += 2 * kMaxVecBlocks * 8;
    double *scal = reinterpret_cast<double *>(w_); w_ += S_COUNT * 8;
    unsigned *ticket = reinterpret_cast<unsigned *>(w_);
    RSG_CUDA_TRY(cudaMemsetAsync(scal, 0, S_COUNT * 8 + 64, st));
    const int grid = std::max(1, std::min(kMaxVecBlocks, std::min(h->sm_count * 8, (n + kVecThreads * 4 - 1) / (kVecThreads * 4))));
    static_assert(kGmresRestart <= 30, "krylov_pin is sized for a restart length of 30");
    double *ph = h->krylov_pin + S_COUNT, *py = ph + R + 2; // grow-once pinned mirror (no allocation per call)

    auto precondition = [&](const double *in, double *out) -> int { // in32 already holds (float) in with fp32 factors
        if (!pre) { RSG_CUDA_TRY(cudaMemcpyAsync(out, in, sizeof(double) * (size_t)n, cudaMemcpyDeviceToDevice, st)); return RSG_OK; }
        if (pre32) {
            RSG_TRY(sptrsv_lu_s(h, n, nnz, base, rowptr, colidx, static_cast<const float *>(lu_values), in32, tmp32, out32));
            widen_kernel<<<grid, kVecThreads, 0, st>>>(n, out32, out);
            h->launches++;
            return RSG_OK;
        }
        return sptrsv_lu_d(h, n, nnz, base, rowptr, colidx, static_cast<const double *>(lu_values), in, t, out);
    };

    double H[R + 1][R], cs[R], sn[R], gv[R + 1], yv[R];
    dot_kernel<<<grid, kVecThreads, 0, st>>>(n, b, b, partials, ticket, scal + S_BB);
    h->launches++;
    double bb = 1.0, stop = 0.0, rr = 0.0;
    int it = 0;
    bool first = true;
    for (;;) {
        // true residual at every (re)start: V_0 = b - A x
        RSG_TRY(spmv_csr_d(h, n, n, nnz, base, rowptr, colidx, values, 1.0, x, 0.0, w));
        residual_kernel<<<grid, kVecThreads, 0, st>>>(n, b, w, V);
        dot_kernel<<<grid, kVecThreads, 0, st>>>(n, V, V, partials, ticket, hcol);
        h->launches += 2;
        RSG_CUDA_TRY(cudaMemcpyAsync(ph, hcol, sizeof(double), cudaMemcpyDeviceToHost, st));
        if (first) RSG_CUDA_TRY(cudaMemcpyAsync(ph + 1, scal + S_BB, sizeof(double), cudaMemcpyDeviceToHost, st));
        RSG_CUDA_TRY(cudaStreamSynchronize(st));
        if (first) { bb = ph[1] > 0 ? ph[1] : 1.0; stop = rtol * rtol * bb; first = false; }
        rr = ph[0];
        if (!(rr > stop) || !std::isfinite(rr) || it >= max_iters) break;
        gmres_scale_kernel<<<grid, kVecThreads, 0, st>>>(n, hcol, V, V, pre32 ? in32 : nullptr);
        h->launches++;
        gv[0] = std::sqrt(rr);
        for (int i = 1; i <= R; ++i) gv[i] = 0;
        int k = 0;
        for (int j = 0; j < R && it < max_iters; ++j) {
            double *vj = V + nd * (size_t)j, *z = Z + nd * (size_t)j;
            RSG_TRY(precondition(vj, z));
            RSG_TRY(spmv_csr_d(h, n, n, nnz, base, rowptr, colidx, values, 1.0, z, 0.0, w));
            for (int i = 0; i <= j; ++i)
                gmres_mgs_kernel<<<grid, kVecThreads, 0, st>>>(n, i ? hcol + i - 1 : nullptr, i ? V + nd * (size_t)(i - 1) : nullptr,
                                                               V + nd * (size_t)i, w, partials, ticket, hcol + i);
            gmres_mgs_kernel<<<grid, kVecThreads, 0, st>>>(n, hcol + j, vj, w, w, partials, ticket, hcol + j + 1);
            h->launches += j + 2;
            RSG_CUDA_TRY(cudaMemcpyAsync(ph, hcol, sizeof(double) * (size_t)(j + 2), cudaMemcpyDeviceToHost, st));
            RSG_CUDA_TRY(cudaStreamSynchronize(st));
            for (int i = 0; i <= j; ++i) H[i][j] = ph[i];
            const double hn = std::sqrt(ph[j + 1]);
            for (int i = 0; i < j; ++i) {
                const double tmp = cs[i] * H[i][j] + sn[i] * H[i + 1][j];
                H[i + 1][j] = -sn[i] * H[i][j] + cs[i] * H[i + 1][j];
                H[i][j] = tmp;
            }
            const double d = std::hypot(H[j][j], hn);
            if (d == 0) break;
            cs[j] = H[j][j] / d;
            sn[j] = hn / d;
            H[j][j] = d;
            gv[j + 1] = -sn[j] * gv[j];
            gv[j] = cs[j] * gv[j];
            ++it;
            k = j + 1;
            rr = gv[j + 1] * gv[j + 1];
            if (!(rr > stop) || !std::isfinite(rr) || hn == 0) break;
            if (j + 1 < R && it < max_iters) {
                gmres_scale_kernel<<<grid, kVecThreads, 0, st>>>(n, hcol + j + 1, w, V + nd * (size_t)(j + 1), pre32 ? in32 : nullptr);
                h->launches++;
            }
        }
        // x += Z y,  H y = g
        for (int i = k - 1; i >= 0; --i) {
            double s = gv[i];
            for (int q = i + 1; q < k; ++q) s -= H[i][q] * yv[q];
            yv[i] = s / H[i][i];
        }
        if (k > 0) {
            for (int i = 0; i < k; ++i) py[i] = yv[i];
            RSG_CUDA_TRY(cudaMemcpyAsync(ycoef, py, sizeof(double) * (size_t)k, cudaMemcpyHostToDevice, st));
            gmres_update_x_kernel<<<grid, kVecThreads, 0, st>>>(n, k, ycoef, Z, nd, x);
            h->launches++;
        }
        if (!(rr > stop) || !std::isfinite(rr) || it >= max_iters || k == 0) break;
    }
    RSG_CUDA_TRY(cudaStreamSynchronize(st));
    RSG_CUDA_TRY(cudaGetLastError());
    if (iters) *iters = it;
    if (relres) *relres = std::sqrt(rr / bb);
    if (!std::isfinite(rr)) return check_err_flag(h) == RSG_DEPENDENCY_TIMEOUT ? RSG_DEPENDENCY_TIMEOUT : RSG_BREAKDOWN; // a hung solve leaves sentinel NaNs
    return RSG_OK;
}
