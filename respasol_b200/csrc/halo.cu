// halo.cu -- x-halo exchange of the row-block partitioned SpMV over NVLink peer memory.
// The reference is single-GPU (new capability); the exchange step the north star names
// is an all-gather of the entries other ranks reference.  Here every rank PUSHES them: one kernel
// reads each boundary entry of the local x once and stores it straight into every peer's halo
// buffer (P2P-mapped symmetric memory, plain st.global over NVLink/NVSwitch), so the step costs one
// small launch plus one cross-GPU barrier instead of an NCCL collective's ~15 us of setup.
#include "rsg_internal.cuh"

namespace rsg {

struct HaloPeers {
    void *dst[RSG_MAX_PEERS]; // peer q's halo slot for THIS rank (already offset), NULL = skip
    int *flag[RSG_MAX_PEERS]; // peer q's arrival flag for THIS rank, NULL = no signal
};

template <typename T>
__global__ void __launch_bounds__(256)
halo_push_kernel(int nsend, const int *__restrict__ send_idx, const T *__restrict__ x_local, int npeers, HaloPeers peers,
                 int epoch, int *cta_counter)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j < nsend) {
        const T v = x_local[send_idx[j]];
#pragma unroll 1
        for (int q = 0; q < npeers; ++q) {
            T *d = static_cast<T *>(peers.dst[q]);
            if (d) d[j] = v;
        }
    }
    if (epoch == 0) return;
    // signal: the last CTA to finish publishes the epoch to every peer once all stores are visible
    // (one thread per peer: the system-scope fences overlap instead of queueing behind each other)
    __shared__ int s_last;
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        const int done = atomicAdd(cta_counter, 1);
        s_last = (done == (int)gridDim.x - 1);
        if (s_last) *cta_counter = 0; // next launch on this stream starts from zero
    }
    __syncthreads();
    if (s_last && threadIdx.x < npeers && peers.flag[threadIdx.x]) {
        __threadfence_system();
        asm volatile("st.release.sys.global.s32 [%0], %1;" ::"l"(peers.flag[threadIdx.x]), "r"(epoch) : "memory");
    }
}

} // namespace rsg

extern "C" int rsg_halo_push(rsg_handle_t h, int elem_bytes, int nsend, const int *send_idx, const void *x_local,
                             int npeers, void *const *peer_slots, int *const *peer_flags, int epoch)
{
    using namespace rsg;
    if (!h || nsend < 0 || npeers < 0 || npeers > RSG_MAX_PEERS || (elem_bytes != 4 && elem_bytes != 8)) return RSG_BAD_ARG;
    if (!peer_flags) epoch = 0;
    if (npeers == 0 || (nsend == 0 && epoch == 0)) return RSG_OK;
    if (nsend > 0 && (!send_idx || !x_local || !peer_slots)) return RSG_BAD_ARG;
    int prev = -1;
    cudaGetDevice(&prev);
    if (prev != h->device) cudaSetDevice(h->device);
    HaloPeers peers;
    for (int q = 0; q < RSG_MAX_PEERS; ++q) {
        peers.dst[q] = (q < npeers && peer_slots) ? peer_slots[q] : nullptr;
        peers.flag[q] = (q < npeers && peer_flags) ? peer_flags[q] : nullptr;
    }
    const int grid = std::max(1, (nsend + 255) / 256);
    int *counter = h->scratch_i + 62; // zero since rsg_create; every signalling launch leaves it at zero
    if (elem_bytes == 8)
        halo_push_kernel<double><<<grid, 256, 0, h->stream>>>(nsend, send_idx, static_cast<const double *>(x_local), npeers, peers,
                                                              epoch, counter);
    else
        halo_push_kernel<float><<<grid, 256, 0, h->stream>>>(nsend, send_idx, static_cast<const float *>(x_local), npeers, peers,
                                                             epoch, counter);
    h->launches++;
    const cudaError_t e = cudaGetLastError();
    if (prev != h->device && prev >= 0) cudaSetDevice(prev);
    return e == cudaSuccess ? RSG_OK : RSG_CUDA_ERROR;
}

extern "C" int rsg_spmv_halo_wait(rsg_handle_t h, const int *flags, int nflags, int epoch)
{
    if (!h || nflags < 0 || nflags > RSG_MAX_PEERS || (nflags > 0 && !flags)) return RSG_BAD_ARG;
    h->wait_flags = nflags > 0 ? flags : nullptr;
    h->wait_n = nflags;
    h->wait_epoch = epoch;
    return RSG_OK;
}

extern "C" int rsg_spmv_halo_fuse_ranges(rsg_handle_t h, int nsend, const int *send_idx, int npeers, void *const *peer_slots,
                                         int *const *peer_flags, const int *peer_lo, const int *peer_hi, int epoch);

extern "C" int rsg_spmv_halo_fuse(rsg_handle_t h, int nsend, const int *send_idx, int npeers, void *const *peer_slots,
                                  int *const *peer_flags, int epoch)
{
    return rsg_spmv_halo_fuse_ranges(h, nsend, send_idx, npeers, peer_slots, peer_flags, nullptr, nullptr, epoch);
}

extern "C" int rsg_spmv_halo_fuse_ranges(rsg_handle_t h, int nsend, const int *send_idx, int npeers, void *const *peer_slots,
                                         int *const *peer_flags, const int *peer_lo, const int *peer_hi, int epoch)
{
    if (!h || nsend < 0 || npeers < 0 || npeers > RSG_MAX_PEERS || epoch <= 0 || !peer_flags) return RSG_BAD_ARG;
    if (nsend > 0 && (!send_idx || !peer_slots)) return RSG_BAD_ARG;
    rsg::HaloFuse &f = h->fuse;
    for (int q = 0; q < RSG_MAX_PEERS; ++q) {
        f.dst[q] = (q < npeers && peer_slots) ? peer_slots[q] : nullptr;
        f.flag[q] = q < npeers ? peer_flags[q] : nullptr;
        f.lo[q] = (q < npeers && peer_lo) ? peer_lo[q] : 0;
        f.hi[q] = (q < npeers && peer_hi) ? peer_hi[q] : nsend;
    }
    f.send_idx = send_idx;
    f.counter = h->scratch_i + 62;
    f.nsend = nsend; f.npeers = npeers; f.epoch = epoch;
    f.nruns = 0;
    f.packets = 0;
    f.pull = 0;
    f.ctas = 1; // sized at launch from the kernel's CTA width
    return RSG_OK;
}
