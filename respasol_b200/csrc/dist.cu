// dist.cu -- the multi-GPU entry points of the C ABI ("multi-GPU entry point takes an array of
// handles"; 8e: row-block SpMV across the box).  No counterpart in the reference, whose harness drives one GPU
// (/root/reference/GPU/spmv.cu:174-207); this is the shape a C++ caller of that harness needs to reach 8 GPUs.
//
//   rsg_dist_*        ONE rank's side of the row-block product, everything a step needs bound once: the block's
//                     CSR with relabelled columns, the two x images ([local | halo], peer-mapped), the peers' halo
//                     slots / arrival flags and the range of the send list each peer reads.  rsg_dist_step = arm the
//                     fused push + the in-kernel wait + launch the product: ONE library call, ONE kernel launch per
//                     step.  Used by the one-process-per-GPU path (respasol_b200/dist.py under torchrun, memory mapped
//                     through torch symmetric memory / CUDA IPC) and by rsg_spmv_multi_* below.
//   rsg_spmv_multi_*  a whole matrix over N devices of ONE process: partition (the rsg_get_row_partition formula, or the
//                     caller's cuts), halo relabelling, peer access (cudaDeviceEnablePeerAccess), one handle + one
//                     rsg_dist per device, one host thread per device issuing its steps.  bin/test_spmv --gpus N.
//
// Three exchange protocols share the two-image scheme below (DESIGN.md 6 has the measurements behind the choice):
//   flags    (rsg_dist_create alone)        push CTAs inside the product's grid, fence, counter, arrival flags; blocks >= 12 M entries
//   packets  (+ rsg_dist_enable_packets)    {value, step} packets from a side-stream kernel, unpacked by the halo tiles; short products
//   pull     (+ rsg_dist_enable_pull)       the product's first CTAs fetch the halo from the owners' x; any pattern, no remote store
// Protocol of the flags form (unchanged from round 1, now in one place): step e uses image p = (e-1) & 1.  A rank's launch first stores
// its boundary entries of image p into the readers' image-p halo slots and releases e to their flags; tiles that read
// halo columns (and, always, the last tile) acquire flags[p][i] >= e from every neighbour.  A rank therefore never runs
// more than one step ahead of a neighbour, which is what makes two images enough.
#include "spmv_device.cuh"

#include <cstdlib>
#include <cstring>
#include <map>
#include <new>
#include <thread>

using namespace rsg;

struct rsg_dist {
    rsg_context *h = nullptr;
    char precision = 'd';
    int m = 0, n_ext = 0, nnz = 0, base = 0, halo_start = 0;
    const int *rowptr = nullptr, *colidx = nullptr;
    const void *values = nullptr;
    void *x_image[2] = {nullptr, nullptr};
    HaloFuse fuse[2];
    const int *wait_flags[2] = {nullptr, nullptr};
    int nwait = 0, epoch = 0;
    long long stored = 0; // entries of the block (from its plan)
    // chained products (rsg_dist_create_chain): three x images, step e reads image e % 3 and writes image (e + 1) % 3
    bool chained = false;
    void *x3[3] = {nullptr, nullptr, nullptr};
    ChainPush cp[3];                 // cp[i]: the push INTO the peers' image i
    const int *wflags3[3] = {nullptr, nullptr, nullptr};
    // packet exchange (rsg_dist_enable_packets): this rank's two packet buffers; fuse[p].dst then point at the peers' ones
    bool packets = false;
    const void *pk_local[2] = {nullptr, nullptr};
    bool pk_side = false;            // the packets are pushed by a small kernel of their own on `side` (no remote store in the product)
    cudaStream_t side = nullptr;
    cudaEvent_t x_ready = nullptr;
    void *slot_dst[2][RSG_MAX_PEERS] = {}; // the peers' plain halo slots (restored when the packets are switched off)
    // pull exchange (rsg_dist_enable_pull): the product's first CTAs fetch the halo; no remote write in the product
    bool pull = false;
    HaloFuse pullf[2];
    const int *local_flag[2] = {nullptr, nullptr};
    int *send_pos = nullptr;         // device, m ints (owned)
    int *chain_counter = nullptr;    // device (owned, with send_pos)
};

namespace {

struct DevGuard {
    int prev = -1;
    explicit DevGuard(int dev) { cudaGetDevice(&prev); if (prev != dev) cudaSetDevice(dev); else prev = -1; }
    ~DevGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

constexpr long long kDistPdlMinEntries = 12000000; // per-rank entries below which a fused-exchange step is launched plainly

// every halo column this block reads: its packet (both images) -> "pending" (epoch field all ones); packets of columns
// nobody reads keep epoch 0 = "padding of an x window, take it as it is"
__global__ void packet_mark_kernel(long long cnt, const int *__restrict__ colidx, int base, int halo_start, void *pk0, void *pk1, int wide)
{
    const long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= cnt) return;
    const int c = colidx[k] - base;
    if (c < halo_start) return;
    const int j = c - halo_start;
    if (wide) {
        static_cast<uint4 *>(pk0)[j] = make_uint4(0u, 0xffffffffu, 0u, 0xffffffffu);
        static_cast<uint4 *>(pk1)[j] = make_uint4(0u, 0xffffffffu, 0u, 0xffffffffu);
    } else {
        static_cast<uint2 *>(pk0)[j] = make_uint2(0u, 0xffffffffu);
        static_cast<uint2 *>(pk1)[j] = make_uint2(0u, 0xffffffffu);
    }
}

// the boundary entries of x as {value, step} packets into the peers' packet buffers (rsg_dist_enable_packets, side-stream form)
template <typename TX> __global__ void __launch_bounds__(128) packet_push_kernel(const HaloFuse fuse, const TX *__restrict__ x)
{
    halo_fuse_push<TX, 128>(fuse, x);
}

inline int halo_base_of(int n_local) { return (n_local + 31) / 32 * 32; } // the halo starts on its own 128/256-byte line

} // namespace

extern "C" {

int rsg_dist_create(rsg_handle_t h, char precision, int m, int n_ext, int nnz, int base, const int *rowptr, const int *colidx,
                    const void *values, void *const *x_images, int halo_start, int nsend, const int *send_idx, int npeers,
                    void *const *peer_slots, int *const *peer_flags, const int *peer_lo, const int *peer_hi,
                    const int *const *wait_flags, int nwait, rsg_dist_t *out)
{
    if (!h || !out || m < 0 || n_ext < 0 || !rowptr || !x_images || !x_images[0] || !x_images[1] || npeers < 0 || npeers > RSG_MAX_PEERS ||
        nwait < 0 || nwait > RSG_MAX_PEERS || (precision != 'd' && precision != 's' && precision != 'm') ||
        (npeers > 0 && (!peer_slots || !peer_flags)) || (nwait > 0 && (!wait_flags || !wait_flags[0] || !wait_flags[1])) ||
        (nsend > 0 && !send_idx))
        return RSG_BAD_ARG;
    *out = nullptr;
    rsg_dist *d = new (std::nothrow) rsg_dist();
    if (!d) return RSG_ALLOC_FAILED;
    d->h = h; d->precision = precision; d->m = m; d->n_ext = n_ext; d->nnz = nnz; d->base = base; d->halo_start = halo_start;
    d->rowptr = rowptr; d->colidx = colidx; d->values = values;
    d->nwait = nwait;
    for (int p = 0; p < 2; ++p) {
        d->x_image[p] = x_images[p];
        d->wait_flags[p] = nwait > 0 ? wait_flags[p] : nullptr;
        HaloFuse &f = d->fuse[p];
        std::memset(&f, 0, sizeof(f));
        for (int q = 0; q < npeers; ++q) { // image-major: [p * npeers + q]
            f.dst[q] = peer_slots[p * npeers + q];
            f.flag[q] = peer_flags[p * npeers + q];
            f.lo[q] = peer_lo ? peer_lo[q] : 0;
            f.hi[q] = peer_hi ? peer_hi[q] : nsend;
        }
        f.send_idx = send_idx;
        f.counter = h->scratch_i + 62;
        f.nsend = nsend; f.npeers = npeers; f.ctas = npeers > 0 ? 1 : 0;
    }
    if (nsend > 0 && npeers > 0) { // a send list made of a few contiguous runs needs no index load in the push
        std::vector<int> idx((size_t)nsend);
        DevGuard g0(h->device);
        if (cudaMemcpy(idx.data(), send_idx, sizeof(int) * (size_t)nsend, cudaMemcpyDeviceToHost) != cudaSuccess) { (void)cudaGetLastError(); delete d; return RSG_CUDA_ERROR; }
        int nruns = 1, start[4] = {idx[0], 0, 0, 0}, len[4] = {1, 0, 0, 0};
        for (int j = 1; j < nsend && nruns <= 4; ++j) {
            if (idx[(size_t)j] == idx[(size_t)j - 1] + 1) { len[nruns - 1]++; continue; }
            if (++nruns > 4) break;
            start[nruns - 1] = idx[(size_t)j]; len[nruns - 1] = 1;
        }
        for (int p = 0; p < 2; ++p) {
            d->fuse[p].nruns = nruns <= 4 ? nruns : 0;
            for (int r = 0; r < 4; ++r) { d->fuse[p].run_start[r] = start[r]; d->fuse[p].run_len[r] = len[r]; }
        }
    }
    // the plan of this block, halo tiles marked, now (so that no step pays for an analysis)
    DevGuard g(h->device);
    const int keep = h->opts[RSG_OPT_SPMV_HALO_START];
    h->opts[RSG_OPT_SPMV_HALO_START] = halo_start;
    SpmvPlan *pl = nullptr;
    const int pkey = precision == 'd' ? 0x88 : precision == 's' ? 0x44 : 0x84;
    const int rc = m > 0 ? spmv_get_plan(h, m, n_ext, nnz, base, rowptr, colidx, false, &pl, pkey) : RSG_OK;
    h->opts[RSG_OPT_SPMV_HALO_START] = keep;
    if (rc != RSG_OK) { delete d; return rc; }
    d->stored = pl ? pl->nnz : 0;
    *out = d;
    return RSG_OK;
}

} // extern "C"

// the (cached) SpMV plan of a dist context's block
static int dist_plan(rsg_dist *d, SpmvPlan **pl)
{
    rsg_context *h = d->h;
    *pl = nullptr;
    if (d->m <= 0) return RSG_OK;
    const int keep = h->opts[RSG_OPT_SPMV_HALO_START];
    h->opts[RSG_OPT_SPMV_HALO_START] = d->halo_start;
    const int pkey = d->precision == 'd' ? 0x88 : d->precision == 's' ? 0x44 : 0x84;
    const int rc = spmv_get_plan(h, d->m, d->n_ext, d->nnz, d->base, d->rowptr, d->colidx, false, pl, pkey);
    h->opts[RSG_OPT_SPMV_HALO_START] = keep;
    return rc;
}

extern "C" {

/* Packet exchange for a dist context made by rsg_dist_create: every boundary entry travels as {value, step number} in one
 * 16-byte (fp32 x: 8-byte) store, and the halo tiles of the product unpack their x window out of the packets, polling the step
 * number inside each entry -- no fence + counter + flag between the data and its consumer (the latency chain that bounds a
 * short product).  local_packets: this rank's 2 packet buffers (image 0, 1; (n_ext - halo_start) packets each, zero-filled by
 * the caller, written by the peers); peer_packets: 2 * npeers pointers, image-major like peer_slots: where entry 0 of this rank's
 * send list lands in peer q's buffer.  ONLY for exchange patterns in which every rank also READS from every rank it sends to
 * (stencils, symmetric patterns): that mutual dependency is what keeps a rank from running two steps ahead of a reader, the
 * job the arrival flags did.  Every rank must have returned from this call before any rank steps (the caller's barrier).
 * RSG_UNSUPPORTED when the block's plan is not the window kernel's. */
int rsg_dist_enable_packets(rsg_dist_t d, void *const *local_packets, void *const *peer_packets, int side_stream)
{
    if (!d || d->chained) return RSG_BAD_ARG;
    rsg_context *h = d->h;
    const int npeers = d->fuse[0].npeers;
    if (!local_packets) { // back to the arrival flags (a caller whose ranks did not all get the window kernel)
        if (d->packets)
            for (int p = 0; p < 2; ++p) {
                for (int q = 0; q < npeers; ++q) d->fuse[p].dst[q] = d->slot_dst[p][q];
                d->fuse[p].packets = 0;
            }
        d->packets = false;
        DevGuard g0(h->device);
        SpmvPlan *pl0 = nullptr;
        RSG_TRY(dist_plan(d, &pl0));
        if (pl0 && pl0->algo == RSG_SPMV_WINDOW) RSG_TRY(window_place_halo(h, *pl0, h->opts[RSG_OPT_SPMV_HALO_PLACE] > 0 ? h->opts[RSG_OPT_SPMV_HALO_PLACE] : 100, 0));
        return RSG_OK;
    }
    if (d->packets || !local_packets[0] || !local_packets[1] || d->halo_start <= 0) return RSG_BAD_ARG;
    for (int p = 0; p < 2; ++p)
        for (int q = 0; q < npeers; ++q)
            if (d->fuse[p].dst[q] && (!peer_packets || !peer_packets[p * npeers + q])) return RSG_BAD_ARG;
    DevGuard g(h->device);
    SpmvPlan *pl = nullptr;
    RSG_TRY(dist_plan(d, &pl));
    if (pl && pl->algo != RSG_SPMV_WINDOW) return RSG_UNSUPPORTED;
    if (pl && pl->nnz > 0) {
        const long long cnt = pl->nnz;
        packet_mark_kernel<<<(unsigned)((cnt + 255) / 256), 256, 0, h->stream>>>(cnt, d->colidx + pl->first, d->base, d->halo_start,
                                                                                  local_packets[0], local_packets[1], d->precision == 's' ? 0 : 1);
        h->launches++;
        RSG_CUDA_TRY(cudaGetLastError());
        RSG_CUDA_TRY(cudaStreamSynchronize(h->stream));
    }
    // halo tiles no longer wait for a flag that trails the data: they can sit in the middle of the grid, and the last wave
    // (the critical path into the next launch) is interior tiles
    // (probe, 8 M entries per rank: 20 % .. 60 % of the grid 15.84-15.87 us, 35 % .. last wave 16.0-16.2, all at the end 20.2)
    if (pl) RSG_TRY(window_place_halo(h, *pl, h->opts[RSG_OPT_SPMV_HALO_PLACE] > 0 ? h->opts[RSG_OPT_SPMV_HALO_PLACE] : 20, 60));
    for (int p = 0; p < 2; ++p) {
        d->pk_local[p] = local_packets[p];
        for (int q = 0; q < npeers; ++q)
            if ((d->slot_dst[p][q] = d->fuse[p].dst[q]) != nullptr) d->fuse[p].dst[q] = peer_packets[p * npeers + q];
        d->fuse[p].packets = 1;
    }
    d->packets = true;
    d->pk_side = false;
    if (side_stream) {
        // The push as a kernel of its own beside the product: a grid that issued remote stores is complete only once they are
        // flushed through NVLink, which delays the launch that depends on it by 2-4.5 us (measured, tools/dist_trace.py); on a
        // side stream that flush is nobody's dependency.  An event record between two products does not cost them their
        // programmatic overlap (measured: 14.15 us with and without).
        int lo_prio = 0, hi_prio = 0;
        cudaDeviceGetStreamPriorityRange(&lo_prio, &hi_prio);
        if (const char *env = getenv("RSG_DIST_SIDE_PRIORITY")) if (!std::strcmp(env, "normal")) hi_prio = 0; // (A/B)
        if (!d->side && cudaStreamCreateWithPriority(&d->side, cudaStreamNonBlocking, hi_prio) != cudaSuccess) { (void)cudaGetLastError(); return RSG_CUDA_ERROR; }
        if (!d->x_ready && cudaEventCreateWithFlags(&d->x_ready, cudaEventDisableTiming) != cudaSuccess) { (void)cudaGetLastError(); return RSG_CUDA_ERROR; }
        d->pk_side = true;
    }
    return RSG_OK;
}

/* Pull exchange for a context made by rsg_dist_create (any pattern, any kernel): instead of pushing its boundary entries, the
 * product's first CTAs FETCH this rank's halo -- they poll the neighbours' "x of step e is final" counters over NVLink, load
 * the entries from the owners' x images and store them into this rank's own image, then raise a local flag the halo tiles
 * wait for.  The product then holds no remote store: nothing to flush through NVLink when the grid completes (which delays the
 * next launch by 2-4 us) and no fence behind remote stores (6-10 us while the memory system is saturated).
 *   npull entries, grouped by owner: pull_col[k] = halo column (in this rank's relabelled numbering) entry k goes to,
 *   pull_idx[k] = its index in the owner's x (DEVICE arrays, caller-owned); owner q's entries are [owner_lo[q], owner_hi[q]).
 *   peer_x: 2 * npeers DEVICE pointers, image-major: peer q's x image p (NULL: nothing is fetched from q).
 *   peer_ready: npeers DEVICE pointers: peer q's counter (NULL: not a neighbour; every neighbour -- source or reader -- is
 *   awaited, so that no rank overwrites an image a reader has not finished with); my_ready: this rank's own counter;
 *   local_flags: 2 ints of this rank (image 0, 1).  All zero-initialised by the caller; every rank enables before any steps. */
int rsg_dist_enable_pull(rsg_dist_t d, int npull, const int *pull_col, const int *pull_idx, int npeers, const void *const *peer_x,
                         const int *const *peer_ready, const int *owner_lo, const int *owner_hi, int *my_ready, int *local_flags)
{
    if (!d || d->chained || d->packets || npull < 0 || npeers < 0 || npeers > RSG_MAX_PEERS || !my_ready || !local_flags ||
        (npull > 0 && (!pull_col || !pull_idx || !peer_x || !owner_lo || !owner_hi)) || (npeers > 0 && !peer_ready) || d->halo_start <= 0)
        return RSG_BAD_ARG;
    rsg_context *h = d->h;
    DevGuard g(h->device);
    SpmvPlan *pl = nullptr;
    RSG_TRY(dist_plan(d, &pl));
    if (pl && pl->algo != RSG_SPMV_WINDOW && pl->algo != RSG_SPMV_TILE) return RSG_UNSUPPORTED;
    // the halo tiles wait for a LOCAL flag that trails the fetch by about two NVLink round trips: mid-grid, last wave interior
    if (pl && pl->algo == RSG_SPMV_WINDOW)
        RSG_TRY(window_place_halo(h, *pl, h->opts[RSG_OPT_SPMV_HALO_PLACE] > 0 ? h->opts[RSG_OPT_SPMV_HALO_PLACE] : 55, 0));
    for (int p = 0; p < 2; ++p) {
        HaloFuse &f = d->pullf[p];
        std::memset(&f, 0, sizeof(f));
        for (int q = 0; q < npeers; ++q) {
            f.dst[q] = peer_x ? const_cast<void *>(peer_x[p * npeers + q]) : nullptr;
            f.flag[q] = const_cast<int *>(peer_ready[q]);
            f.lo[q] = owner_lo ? owner_lo[q] : 0;
            f.hi[q] = owner_hi ? owner_hi[q] : 0;
        }
        f.send_idx = pull_idx;
        f.pull_col = pull_col;
        f.counter = h->scratch_i + 62;
        f.nsend = npull; f.npeers = npeers; f.ctas = 1;
        f.pull = 1;
        f.local_x = d->x_image[p];
        f.my_ready = my_ready;
        f.local_flag = local_flags + p;
        d->local_flag[p] = local_flags + p;
    }
    d->pull = true;
    return RSG_OK;
}

int rsg_dist_destroy(rsg_dist_t d)
{
    if (!d) return RSG_BAD_ARG;
    if (d->send_pos) { DevGuard g(d->h->device); cudaFree(d->send_pos); }
    if (d->side || d->x_ready) {
        DevGuard g(d->h->device);
        if (d->side) { cudaStreamSynchronize(d->side); cudaStreamDestroy(d->side); }
        if (d->x_ready) cudaEventDestroy(d->x_ready);
    }
    delete d;
    return RSG_OK;
}

void *rsg_dist_x(rsg_dist_t d) // the image the NEXT step reads
{
    if (!d) return nullptr;
    return d->chained ? d->x3[(d->epoch + 1) % 3] : d->x_image[d->epoch & 1];
}

int rsg_dist_step(rsg_dist_t d, void *y)
{
    if (!d || (d->m > 0 && !y)) return RSG_BAD_ARG;
    rsg_context *h = d->h;
    DevGuard g(h->device);
    if (d->epoch >= 0x7ffffff0) return RSG_UNSUPPORTED; // step numbers are compared as ints (and 0 / -1 mean "padding" / "pending" in a packet)
    const int p = d->epoch & 1; // step e = epoch + 1 uses image (e - 1) & 1
    d->epoch += 1;
    // timing experiments only (tools/dist_trace.py; the results are wrong): RSG_OPT_SPMV_X_HINT bit 8 = nobody waits for / unpacks
    // the halo, bit 16 = no push (implies bit 8), bit 32 = never fall back to plain launches, bit 64 = an eighth of the push CTAs
    int dbg = h->opts[RSG_OPT_SPMV_X_HINT] & (8 | 16 | 32 | 64 | 128);
    if (dbg & 16) dbg |= 8; // (waiting for a push that never comes would hang until the spin limit)
    if (d->pull) {
        if (!(dbg & 16)) { h->fuse = d->pullf[p]; h->fuse.epoch = d->epoch; }
    } else if (d->packets && d->pk_side) {
        if (d->fuse[p].npeers > 0 && d->fuse[p].nsend > 0 && !(dbg & 16)) {
            // x of this step = everything queued on the product's stream so far; the push follows it on the side stream
            RSG_CUDA_TRY(cudaEventRecord(d->x_ready, h->stream));
            RSG_CUDA_TRY(cudaStreamWaitEvent(d->side, d->x_ready, 0));
            HaloFuse f = d->fuse[p];
            f.epoch = d->epoch;
            f.ctas = std::max(1, std::min(1024, (f.nsend + 4 * 128 - 1) / (4 * 128)));
            if (dbg & 128) f.ctas = std::max(1, std::min(1024, (f.nsend + 127) / 128)); // (experiment: one entry per thread)
            if (dbg & 64) f.ctas = std::max(1, f.ctas / 8); // (experiment: few CTAs, each walking eight batches)
            if (d->precision == 's') packet_push_kernel<float><<<f.ctas, 128, 0, d->side>>>(f, static_cast<const float *>(d->x_image[p]));
            else packet_push_kernel<double><<<f.ctas, 128, 0, d->side>>>(f, static_cast<const double *>(d->x_image[p]));
            h->launches++;
            RSG_CUDA_TRY(cudaGetLastError());
        }
    } else if (d->fuse[p].npeers > 0 && !(dbg & 16)) {
        h->fuse = d->fuse[p];
        h->fuse.epoch = d->epoch;
    }
    if (dbg & 8) { /* no wait */ }
    else if (d->pull) { h->wait_flags = d->local_flag[p]; h->wait_n = 1; h->wait_epoch = d->epoch; }
    else if (d->packets) { h->packets.pkts = d->pk_local[p]; h->packets.halo_start = d->halo_start; h->wait_epoch = d->epoch; }
    else if (d->nwait > 0) { h->wait_flags = d->wait_flags[p]; h->wait_n = d->nwait; h->wait_epoch = d->epoch; }
    const int keep = h->opts[RSG_OPT_SPMV_HALO_START], keep_pdl = h->opts[RSG_OPT_SPMV_PDL];
    h->opts[RSG_OPT_SPMV_HALO_START] = d->halo_start; // part of the plan's key
    // A step cannot be shorter than push + signal + the neighbours' halo tiles (the push of step e+1 follows the end of step e,
    // which follows the arrival of the neighbours' push e): about 15 us while a dependent-launched product saturates the
    // memory system (every hop of the push queues behind its burst), about 1 us more than the product itself with plain
    // launches (profiles/r2_dist_probe.jsonl).  Short per-rank products therefore keep plain launches.
    // (The packet exchange has no such chain: its products keep the dependent launches.)
    if (!d->packets && !d->pull && !(dbg & 32) && d->fuse[p].npeers > 0 && d->stored < kDistPdlMinEntries && keep_pdl != 1) h->opts[RSG_OPT_SPMV_PDL] = 1;
    int rc;
    if (d->precision == 'd')
        rc = spmv_run<double, double>(h, d->m, d->n_ext, d->nnz, d->base, d->rowptr, d->colidx, static_cast<const double *>(d->values), 1.0,
                                      static_cast<const double *>(d->x_image[p]), 0.0, static_cast<double *>(y));
    else if (d->precision == 's')
        rc = spmv_run<float, float>(h, d->m, d->n_ext, d->nnz, d->base, d->rowptr, d->colidx, static_cast<const float *>(d->values), 1.0f,
                                    static_cast<const float *>(d->x_image[p]), 0.0f, static_cast<float *>(y));
    else
        rc = spmv_run<float, double>(h, d->m, d->n_ext, d->nnz, d->base, d->rowptr, d->colidx, static_cast<const float *>(d->values), 1.0,
                                     static_cast<const double *>(d->x_image[p]), 0.0, static_cast<double *>(y));
    h->opts[RSG_OPT_SPMV_HALO_START] = keep;
    h->opts[RSG_OPT_SPMV_PDL] = keep_pdl;
    h->wait_flags = nullptr;
    h->packets.pkts = nullptr;
    h->fuse.ctas = 0;
    return rc;
}

/* Chained products x_{k+1} = alpha * A * x_k on one rank's block (SURVEY.md 8e's use case: "repeated products, x_{k+1} <- y_k").
 * Three x images of the [local | halo] layout rotate: step e reads image e % 3 and writes its rows of the product into the local
 * part of image (e + 1) % 3; the tiles that own rows a neighbour reads also store them into that neighbour's halo slots of
 * image (e + 1) % 3 and, when the last of them is done, release e to the neighbour's flag of that image.  Step e + 1 acquires
 * flags >= e before its halo tiles copy x.  The exchange of a step thus has the whole product to complete in, instead of
 * sitting between two products.  peer_slots / peer_flags: 3 * npeers pointers, image-major; wait_flags: 3 pointers.
 * Before the first step the caller fills image 1 completely (local entries AND halo; rsg_dist_x gives the pointer).
 * Needs the window kernel (RSG_UNSUPPORTED for patterns it declines). */
int rsg_dist_create_chain(rsg_handle_t h, char precision, int m, int n_ext, int nnz, int base, const int *rowptr, const int *colidx,
                          const void *values, void *const *x_images, int halo_start, int nsend, const int *send_idx, int npeers,
                          void *const *peer_slots, int *const *peer_flags, const int *peer_lo, const int *peer_hi,
                          const int *const *wait_flags, int nwait, rsg_dist_t *out)
{
    if (!h || !out || m <= 0 || n_ext < m || !rowptr || !x_images || !x_images[0] || !x_images[1] || !x_images[2] || npeers < 0 ||
        npeers > RSG_MAX_PEERS || nwait < 0 || nwait > RSG_MAX_PEERS || (precision != 'd' && precision != 's' && precision != 'm') ||
        (npeers > 0 && (!peer_slots || !peer_flags)) || (nwait > 0 && (!wait_flags || !wait_flags[0] || !wait_flags[1] || !wait_flags[2])) ||
        (nsend > 0 && !send_idx))
        return RSG_BAD_ARG;
    *out = nullptr;
    rsg_dist *d = new (std::nothrow) rsg_dist();
    if (!d) return RSG_ALLOC_FAILED;
    d->h = h; d->precision = precision; d->m = m; d->n_ext = n_ext; d->nnz = nnz; d->base = base; d->halo_start = halo_start;
    d->rowptr = rowptr; d->colidx = colidx; d->values = values;
    d->nwait = nwait; d->chained = true;
    if (npeers > 0 && nsend == 0) { delete d; return RSG_UNSUPPORTED; } // a rank nobody reads from has no producer tile to signal with: use rsg_dist_create
    DevGuard g(h->device);
    // send_pos[row] = position of the row in the send list (-1: nobody reads it)
    std::vector<int> idx((size_t)(nsend > 0 ? nsend : 1)), pos((size_t)m + 1, -1);
    if (nsend > 0 && cudaMemcpy(idx.data(), send_idx, sizeof(int) * (size_t)nsend, cudaMemcpyDeviceToHost) != cudaSuccess) { (void)cudaGetLastError(); delete d; return RSG_CUDA_ERROR; }
    for (int j = 0; j < nsend; ++j) {
        if (idx[(size_t)j] < 0 || idx[(size_t)j] >= m) { delete d; return RSG_BAD_ARG; }
        pos[(size_t)idx[(size_t)j]] = j;
    }
    pos[(size_t)m] = 0; // slot m = the counter of finished producer tiles
    if (cudaMalloc(&d->send_pos, sizeof(int) * ((size_t)m + 1)) != cudaSuccess ||
        cudaMemcpy(d->send_pos, pos.data(), sizeof(int) * ((size_t)m + 1), cudaMemcpyHostToDevice) != cudaSuccess) {
        (void)cudaGetLastError();
        if (d->send_pos) cudaFree(d->send_pos);
        delete d;
        return RSG_ALLOC_FAILED;
    }
    d->chain_counter = d->send_pos + m;
    for (int i = 0; i < 3; ++i) {
        d->x3[i] = x_images[i];
        d->wflags3[i] = nwait > 0 ? wait_flags[i] : nullptr;
        ChainPush &c = d->cp[i];
        std::memset(&c, 0, sizeof(c));
        for (int q = 0; q < npeers; ++q) {
            c.dst[q] = peer_slots[i * npeers + q];
            c.flag[q] = peer_flags[i * npeers + q];
            c.lo[q] = peer_lo ? peer_lo[q] : 0;
            c.hi[q] = peer_hi ? peer_hi[q] : nsend;
        }
        c.send_pos = d->send_pos;
        c.counter = d->chain_counter;
        c.npeers = npeers;
    }
    const int keep = h->opts[RSG_OPT_SPMV_HALO_START];
    h->opts[RSG_OPT_SPMV_HALO_START] = halo_start;
    h->plan_send_pos = npeers > 0 ? d->send_pos : nullptr;
    SpmvPlan *pl = nullptr;
    const int pkey = precision == 'd' ? 0x88 : precision == 's' ? 0x44 : 0x84;
    int rc = spmv_get_plan(h, m, n_ext, nnz, base, rowptr, colidx, false, &pl, pkey);
    h->opts[RSG_OPT_SPMV_HALO_START] = keep;
    h->plan_send_pos = nullptr;
    if (rc == RSG_OK && npeers > 0 && (pl->algo != RSG_SPMV_WINDOW || (nsend > 0 && pl->wprod <= 0))) rc = RSG_UNSUPPORTED;
    if (rc != RSG_OK) { cudaFree(d->send_pos); delete d; return rc; }
    d->stored = pl->nnz;
    *out = d;
    return RSG_OK;
}

/* restart the chain (after the caller has refilled image 1 and zeroed the flags): the next step is step 1 */
int rsg_dist_chain_reset(rsg_dist_t d)
{
    if (!d || !d->chained) return RSG_BAD_ARG;
    d->epoch = 0;
    return RSG_OK;
}

int rsg_dist_step_chain(rsg_dist_t d, double alpha)
{
    if (!d || !d->chained) return RSG_BAD_ARG;
    rsg_context *h = d->h;
    DevGuard g(h->device);
    const int e = ++d->epoch;             // this step
    const int src = e % 3, dstimg = (e + 1) % 3;
    if (d->cp[dstimg].npeers > 0) {
        h->chain = d->cp[dstimg];
        h->chain.epoch = e;
    }
    if (d->nwait > 0) { h->wait_flags = d->wflags3[src]; h->wait_n = d->nwait; h->wait_epoch = e - 1; }
    const int keep = h->opts[RSG_OPT_SPMV_HALO_START];
    h->opts[RSG_OPT_SPMV_HALO_START] = d->halo_start;
    h->plan_send_pos = d->cp[dstimg].npeers > 0 ? d->send_pos : nullptr; // part of the plan's key
    int rc;
    if (d->precision == 'd')
        rc = spmv_run<double, double>(h, d->m, d->n_ext, d->nnz, d->base, d->rowptr, d->colidx, static_cast<const double *>(d->values), alpha,
                                      static_cast<const double *>(d->x3[src]), 0.0, static_cast<double *>(d->x3[dstimg]));
    else if (d->precision == 's')
        rc = spmv_run<float, float>(h, d->m, d->n_ext, d->nnz, d->base, d->rowptr, d->colidx, static_cast<const float *>(d->values), (float)alpha,
                                    static_cast<const float *>(d->x3[src]), 0.0f, static_cast<float *>(d->x3[dstimg]));
    else
        rc = spmv_run<float, double>(h, d->m, d->n_ext, d->nnz, d->base, d->rowptr, d->colidx, static_cast<const float *>(d->values), alpha,
                                     static_cast<const double *>(d->x3[src]), 0.0, static_cast<double *>(d->x3[dstimg]));
    h->opts[RSG_OPT_SPMV_HALO_START] = keep;
    h->plan_send_pos = nullptr;
    h->wait_flags = nullptr;
    h->chain.npeers = 0;
    return rc;
}

} // extern "C"

// ---------------------------------------------------------------------------------------------------------------
// Single process, N devices.
struct rsg_multi {
    int P = 0;
    char precision = 'd';
    int m = 0, n = 0, base = 0;
    size_t vs = 8, xs = 8;
    std::vector<int> dev, row_start, n_local, hbase;
    int width = 1;
    size_t stride = 0; // elements per x image
    std::vector<rsg_context *> h;
    std::vector<rsg_dist *> d, dc; // separate products / chained products (dc[r] NULL when the pattern has no chained form)
    std::vector<std::vector<int>> halo_src; // per rank: global column behind every halo slot entry (-1 unused)
    std::vector<cudaStream_t> stream;
    std::vector<cudaEvent_t> ev0, ev1;
    // per device allocations
    std::vector<int *> rowptr, colidx, send_idx, flags;
    std::vector<void *> values, x, y, pk; // pk: packet buffers (2 images) when the packet exchange is on
    bool packets = false;
    int proto = 0; // 0 arrival flags, 1 packets, 2 pull
    std::vector<int *> pull_col, pull_idx;
    long long bytes_received_per_step = 0; // sum over ranks of the halo entries they read, in bytes
    long long halo_tiles = 0;
};

namespace {

void multi_free(rsg_multi *M)
{
    for (int r = 0; r < M->P; ++r) {
        if ((size_t)r >= M->dev.size()) break;
        DevGuard g(M->dev[r]);
        if ((size_t)r < M->stream.size() && M->stream[r]) cudaStreamSynchronize(M->stream[r]);
        if ((size_t)r < M->d.size() && M->d[r]) rsg_dist_destroy(M->d[r]);
        if ((size_t)r < M->dc.size() && M->dc[r]) rsg_dist_destroy(M->dc[r]);
        if ((size_t)r < M->h.size() && M->h[r]) rsg_destroy(M->h[r]);
        if ((size_t)r < M->rowptr.size()) cudaFree(M->rowptr[r]);
        if ((size_t)r < M->colidx.size()) cudaFree(M->colidx[r]);
        if ((size_t)r < M->send_idx.size()) cudaFree(M->send_idx[r]);
        if ((size_t)r < M->flags.size()) cudaFree(M->flags[r]);
        if ((size_t)r < M->values.size()) cudaFree(M->values[r]);
        if ((size_t)r < M->x.size()) cudaFree(M->x[r]);
        if ((size_t)r < M->y.size()) cudaFree(M->y[r]);
        if ((size_t)r < M->pk.size() && M->pk[r]) cudaFree(M->pk[r]);
        if ((size_t)r < M->pull_col.size() && M->pull_col[r]) cudaFree(M->pull_col[r]);
        if ((size_t)r < M->pull_idx.size() && M->pull_idx[r]) cudaFree(M->pull_idx[r]);
        if ((size_t)r < M->ev0.size() && M->ev0[r]) cudaEventDestroy(M->ev0[r]);
        if ((size_t)r < M->ev1.size() && M->ev1[r]) cudaEventDestroy(M->ev1[r]);
        if ((size_t)r < M->stream.size() && M->stream[r]) cudaStreamDestroy(M->stream[r]);
    }
    delete M;
}

// one host thread per device: what a C++ caller with N GPUs does anyway, and the only way N launch queues fill at once
template <typename F> int for_each_rank(rsg_multi *M, F fn)
{
    std::vector<int> rc((size_t)M->P, RSG_OK);
    if (M->P == 1) { rc[0] = fn(0); return rc[0]; }
    std::vector<std::thread> th;
    th.reserve((size_t)M->P);
    for (int r = 0; r < M->P; ++r) th.emplace_back([&, r]() { cudaSetDevice(M->dev[r]); rc[(size_t)r] = fn(r); });
    for (auto &t : th) t.join();
    for (int r = 0; r < M->P; ++r)
        if (rc[(size_t)r] != RSG_OK) return rc[(size_t)r];
    return RSG_OK;
}

} // namespace

extern "C" {

int rsg_spmv_multi_create(int ngpu, const int *devices, char precision, int m, int n, int nnz, int base, const int *rowptr,
                          const int *colidx, const double *values, const int *row_starts, rsg_multi_t *out)
{
    (void)nnz;
    if (!out || ngpu < 1 || ngpu > RSG_MAX_PEERS || m < 0 || n < 0 || m != n || !rowptr || (base != 0 && base != 1) ||
        (precision != 'd' && precision != 's' && precision != 'm'))
        return RSG_BAD_ARG; // square matrices: x is distributed like the rows (the reference harness assumes square too, spmv.cu:50-51)
    *out = nullptr;
    int count = 0;
    RSG_CUDA_TRY(cudaGetDeviceCount(&count));
    rsg_multi *M = new (std::nothrow) rsg_multi();
    if (!M) return RSG_ALLOC_FAILED;
    const int P = ngpu;
    M->P = P; M->precision = precision; M->m = m; M->n = n; M->base = base;
    M->vs = precision == 'd' ? 8 : 4; M->xs = precision == 's' ? 4 : 8;
    M->dev.resize((size_t)P);
    for (int r = 0; r < P; ++r) {
        M->dev[(size_t)r] = devices ? devices[r] : r;
        if (M->dev[(size_t)r] < 0 || M->dev[(size_t)r] >= count) { delete M; return RSG_BAD_ARG; }
    }
    // ---- partition: the caller's cuts, or nnz-balanced split points (rsg_get_row_partition's formula, 64-bit)
    std::vector<int> &rs = M->row_start;
    rs.assign((size_t)P + 1, 0);
    const long long first = rowptr[0], total = (long long)rowptr[m] - first;
    for (int g = 1; g < P; ++g) {
        if (row_starts) rs[(size_t)g] = row_starts[g];
        else {
            const long long target = (long long)g * total / P + first;
            rs[(size_t)g] = (int)(std::lower_bound(rowptr, rowptr + m + 1, target, [](int v, long long t) { return (long long)v < t; }) - rowptr);
        }
        if (rs[(size_t)g] < rs[(size_t)g - 1] || rs[(size_t)g] > m) { delete M; return RSG_BAD_ARG; }
    }
    rs[(size_t)P] = m;
    for (int g = 0; g < P && P > 1; ++g)
        if (rs[(size_t)g + 1] == rs[(size_t)g]) { delete M; return RSG_BAD_ARG; } // a rank without rows would never signal its neighbours
    auto owner = [&](int c) { return (int)(std::upper_bound(rs.begin(), rs.end(), c) - rs.begin()) - 1; };
    // ---- what every rank reads from the others, what every owner publishes
    std::vector<std::vector<int>> need((size_t)P), send((size_t)P);
    for (int r = 0; r < P; ++r) {
        const int lo = rs[(size_t)r], hi = rs[(size_t)r + 1];
        std::vector<int> &nd = need[(size_t)r];
        for (long long k = rowptr[lo] - base; k < rowptr[hi] - base; ++k) {
            const int c = colidx[k] - base;
            if (c < 0 || c >= n) { delete M; return RSG_BAD_ARG; }
            if (c < lo || c >= hi) nd.push_back(c);
        }
        std::sort(nd.begin(), nd.end());
        nd.erase(std::unique(nd.begin(), nd.end()), nd.end());
    }
    for (int o = 0; o < P; ++o) {
        std::vector<int> &s = send[(size_t)o];
        for (int q = 0; q < P; ++q)
            if (q != o) for (int c : need[(size_t)q]) if (c >= rs[(size_t)o] && c < rs[(size_t)o + 1]) s.push_back(c - rs[(size_t)o]);
        std::sort(s.begin(), s.end());
        s.erase(std::unique(s.begin(), s.end()), s.end());
        M->width = std::max(M->width, (int)s.size());
    }
    const int width = M->width;
    M->n_local.resize((size_t)P); M->hbase.resize((size_t)P);
    size_t stride = 0;
    for (int r = 0; r < P; ++r) {
        M->n_local[(size_t)r] = rs[(size_t)r + 1] - rs[(size_t)r];
        M->hbase[(size_t)r] = halo_base_of(M->n_local[(size_t)r]);
        stride = std::max(stride, (size_t)M->hbase[(size_t)r] + (size_t)P * (size_t)width);
        M->bytes_received_per_step += (long long)need[(size_t)r].size() * (long long)M->xs;
    }
    stride = (stride + 31) / 32 * 32;
    M->stride = stride;
    // reads[q] = owners q reads from; neighbours[q] = ranks q exchanges entries with, in either direction (sorted)
    std::vector<std::vector<int>> reads((size_t)P), nb((size_t)P);
    for (int q = 0; q < P; ++q) {
        std::vector<char> hit((size_t)P, 0);
        for (int c : need[(size_t)q]) hit[(size_t)owner(c)] = 1;
        for (int o = 0; o < P; ++o) if (hit[(size_t)o] && o != q) reads[(size_t)q].push_back(o);
    }
    for (int q = 0; q < P; ++q) {
        std::vector<char> is((size_t)P, 0);
        for (int o : reads[(size_t)q]) is[(size_t)o] = 1;
        for (int o = 0; o < P; ++o) for (int t : reads[(size_t)o]) if (t == q) is[(size_t)o] = 1;
        for (int o = 0; o < P; ++o) if (is[(size_t)o]) nb[(size_t)q].push_back(o);
    }
    // ---- devices: peer access, handles, arrays
    M->h.assign((size_t)P, nullptr); M->d.assign((size_t)P, nullptr); M->dc.assign((size_t)P, nullptr); M->halo_src.assign((size_t)P, std::vector<int>()); M->stream.assign((size_t)P, nullptr);
    M->ev0.assign((size_t)P, nullptr); M->ev1.assign((size_t)P, nullptr);
    M->rowptr.assign((size_t)P, nullptr); M->colidx.assign((size_t)P, nullptr); M->send_idx.assign((size_t)P, nullptr);
    M->flags.assign((size_t)P, nullptr); M->values.assign((size_t)P, nullptr); M->x.assign((size_t)P, nullptr); M->y.assign((size_t)P, nullptr);
    int prev = 0;
    cudaGetDevice(&prev);
    int rc = RSG_OK;
    for (int r = 0; r < P && rc == RSG_OK; ++r) {
        cudaSetDevice(M->dev[(size_t)r]);
        for (int q = 0; q < P; ++q) {
            if (q == r || M->dev[(size_t)q] == M->dev[(size_t)r]) continue;
            int can = 0;
            cudaDeviceCanAccessPeer(&can, M->dev[(size_t)r], M->dev[(size_t)q]);
            if (!can) { rc = RSG_UNSUPPORTED; break; }
            const cudaError_t e = cudaDeviceEnablePeerAccess(M->dev[(size_t)q], 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) { rc = RSG_CUDA_ERROR; break; }
            (void)cudaGetLastError();
        }
        if (rc != RSG_OK) break;
        rc = rsg_create(&M->h[(size_t)r], M->dev[(size_t)r]);
        if (rc != RSG_OK) break;
        if (cudaStreamCreateWithFlags(&M->stream[(size_t)r], cudaStreamNonBlocking) != cudaSuccess ||
            cudaEventCreate(&M->ev0[(size_t)r]) != cudaSuccess || cudaEventCreate(&M->ev1[(size_t)r]) != cudaSuccess) { rc = RSG_CUDA_ERROR; break; }
        rsg_set_stream(M->h[(size_t)r], M->stream[(size_t)r]);
        const int lo = rs[(size_t)r], hi = rs[(size_t)r + 1], ml = hi - lo;
        const long long k0 = rowptr[lo] - base, k1 = rowptr[hi] - base;
        const size_t cnt = (size_t)(k1 - k0);
        // relabel: own columns -> local index; others -> halo_base + owner * width + position in the owner's send list
        std::vector<int> rp((size_t)ml + 1), ci(cnt ? cnt : 1);
        for (int i = 0; i <= ml; ++i) rp[(size_t)i] = (int)((long long)rowptr[lo + i] - base - k0);
        for (size_t k = 0; k < cnt; ++k) {
            const int c = colidx[k0 + (long long)k] - base;
            if (c >= lo && c < hi) ci[k] = c - lo;
            else {
                const int o = owner(c);
                const std::vector<int> &s = send[(size_t)o];
                const int pos = (int)(std::lower_bound(s.begin(), s.end(), c - rs[(size_t)o]) - s.begin());
                ci[k] = M->hbase[(size_t)r] + o * width + pos;
            }
        }
        bool ok = cudaMalloc(&M->rowptr[(size_t)r], sizeof(int) * ((size_t)ml + 1)) == cudaSuccess &&
                  cudaMalloc(&M->colidx[(size_t)r], sizeof(int) * (cnt ? cnt : 1)) == cudaSuccess &&
                  cudaMalloc(&M->values[(size_t)r], M->vs * (cnt ? cnt : 1)) == cudaSuccess &&
                  cudaMalloc(&M->x[(size_t)r], M->xs * 3 * stride) == cudaSuccess &&
                  cudaMalloc(&M->y[(size_t)r], M->xs * (size_t)(ml ? ml : 1)) == cudaSuccess &&
                  cudaMalloc(&M->flags[(size_t)r], sizeof(int) * 192) == cudaSuccess &&
                  cudaMalloc(&M->send_idx[(size_t)r], sizeof(int) * (send[(size_t)r].size() ? send[(size_t)r].size() : 1)) == cudaSuccess;
        if (!ok) { (void)cudaGetLastError(); rc = RSG_ALLOC_FAILED; break; }
        ok = cudaMemcpy(M->rowptr[(size_t)r], rp.data(), sizeof(int) * ((size_t)ml + 1), cudaMemcpyHostToDevice) == cudaSuccess &&
             (cnt == 0 || cudaMemcpy(M->colidx[(size_t)r], ci.data(), sizeof(int) * cnt, cudaMemcpyHostToDevice) == cudaSuccess) &&
             cudaMemset(M->x[(size_t)r], 0, M->xs * 3 * stride) == cudaSuccess && cudaMemset(M->flags[(size_t)r], 0, sizeof(int) * 192) == cudaSuccess &&
             (send[(size_t)r].empty() || cudaMemcpy(M->send_idx[(size_t)r], send[(size_t)r].data(), sizeof(int) * send[(size_t)r].size(), cudaMemcpyHostToDevice) == cudaSuccess);
        if (ok && cnt) {
            if (M->vs == 8) ok = cudaMemcpy(M->values[(size_t)r], values + k0, 8 * cnt, cudaMemcpyHostToDevice) == cudaSuccess;
            else { // fp32 storage: cast on the host like the harness (spmv.cu:63-67)
                std::vector<float> vf(cnt);
                for (size_t k = 0; k < cnt; ++k) vf[k] = (float)values[k0 + (long long)k];
                ok = cudaMemcpy(M->values[(size_t)r], vf.data(), 4 * cnt, cudaMemcpyHostToDevice) == cudaSuccess;
            }
        }
        if (!ok) { (void)cudaGetLastError(); rc = RSG_CUDA_ERROR; break; }
    }
    // ---- per rank: the peers' slots / flags, the send-list range each peer reads, the dist context
    for (int r = 0; r < P && rc == RSG_OK; ++r) {
        std::vector<void *> slots((size_t)(2 * P), nullptr), slots3((size_t)(3 * P), nullptr);
        std::vector<int *> pflags((size_t)(2 * P), nullptr), pflags3((size_t)(3 * P), nullptr);
        std::vector<int> plo((size_t)P, 0), phi((size_t)P, 0);
        const std::vector<int> &s = send[(size_t)r];
        for (int q = 0; q < P; ++q) {
            if (q == r) continue;
            const bool q_reads_r = std::find(reads[(size_t)q].begin(), reads[(size_t)q].end(), r) != reads[(size_t)q].end();
            const auto it = std::find(nb[(size_t)q].begin(), nb[(size_t)q].end(), r);
            if (q_reads_r) { // bounding range, inside r's send list, of what q needs from r
                int lo = (int)s.size(), hi = 0;
                for (int c : need[(size_t)q])
                    if (c >= rs[(size_t)r] && c < rs[(size_t)r + 1]) {
                        const int pos = (int)(std::lower_bound(s.begin(), s.end(), c - rs[(size_t)r]) - s.begin());
                        lo = std::min(lo, pos); hi = std::max(hi, pos + 1);
                    }
                plo[(size_t)q] = lo; phi[(size_t)q] = hi;
            }
            for (int p = 0; p < 2; ++p) {
                if (q_reads_r)
                    slots[(size_t)(p * P + q)] = static_cast<char *>(M->x[(size_t)q]) + M->xs * ((size_t)p * stride + (size_t)M->hbase[(size_t)q] + (size_t)r * (size_t)width);
                if (it != nb[(size_t)q].end())
                    pflags[(size_t)(p * P + q)] = M->flags[(size_t)q] + 32 * p + (int)(it - nb[(size_t)q].begin());
            }
            for (int p = 0; p < 3; ++p) { // chained products: three images, their flags behind the two-image ones
                if (q_reads_r)
                    slots3[(size_t)(p * P + q)] = static_cast<char *>(M->x[(size_t)q]) + M->xs * ((size_t)p * stride + (size_t)M->hbase[(size_t)q] + (size_t)r * (size_t)width);
                if (it != nb[(size_t)q].end())
                    pflags3[(size_t)(p * P + q)] = M->flags[(size_t)q] + 64 + 32 * p + (int)(it - nb[(size_t)q].begin());
            }
        }
        void *ximg[2] = {M->x[(size_t)r], static_cast<char *>(M->x[(size_t)r]) + M->xs * stride};
        const int *wf[2] = {M->flags[(size_t)r], M->flags[(size_t)r] + 32};
        const int ml = M->n_local[(size_t)r];
        cudaSetDevice(M->dev[(size_t)r]);
        rc = rsg_dist_create(M->h[(size_t)r], precision, ml, M->hbase[(size_t)r] + P * width, 0, 0, M->rowptr[(size_t)r], M->colidx[(size_t)r],
                             M->values[(size_t)r], ximg, P > 1 ? M->hbase[(size_t)r] : 0, (int)s.size(), M->send_idx[(size_t)r], P > 1 ? P : 0,
                             slots.data(), pflags.data(), plo.data(), phi.data(), wf, P > 1 ? (int)nb[(size_t)r].size() : 0, &M->d[(size_t)r]);
        if (rc != RSG_OK) break;
        void *ximg3[3] = {M->x[(size_t)r], static_cast<char *>(M->x[(size_t)r]) + M->xs * stride, static_cast<char *>(M->x[(size_t)r]) + M->xs * 2 * stride};
        const int *wf3[3] = {M->flags[(size_t)r] + 64, M->flags[(size_t)r] + 96, M->flags[(size_t)r] + 128};
        const int rcc = rsg_dist_create_chain(M->h[(size_t)r], precision, ml, M->hbase[(size_t)r] + P * width, 0, 0, M->rowptr[(size_t)r],
                                              M->colidx[(size_t)r], M->values[(size_t)r], ximg3, P > 1 ? M->hbase[(size_t)r] : 0, (int)s.size(),
                                              M->send_idx[(size_t)r], P > 1 ? P : 0, slots3.data(), pflags3.data(), plo.data(), phi.data(), wf3,
                                              P > 1 ? (int)nb[(size_t)r].size() : 0, &M->dc[(size_t)r]);
        if (rcc != RSG_OK && rcc != RSG_UNSUPPORTED) rc = rcc; // (unsupported: rsg_spmv_multi_chain will say so)
        // the global column behind every entry of this rank's halo (for the initial fill of a chain)
        std::vector<int> &hs = M->halo_src[(size_t)r];
        hs.assign((size_t)P * (size_t)width, -1);
        for (int o = 0; o < P; ++o) {
            if (o == r) continue;
            const std::vector<int> &so = send[(size_t)o];
            for (int c : need[(size_t)r])
                if (c >= rs[(size_t)o] && c < rs[(size_t)o + 1]) {
                    const int pos = (int)(std::lower_bound(so.begin(), so.end(), c - rs[(size_t)o]) - so.begin());
                    hs[(size_t)o * (size_t)width + (size_t)pos] = c;
                }
        }
    }
    // ---- packet exchange (rsg_dist_enable_packets) when every rank reads from every rank it sends to and every block runs
    // the window kernel; RSG_DIST_PACKETS=0 in the environment keeps the arrival flags (A/B)
    // RSG_DIST_EXCHANGE in the environment: "packets" (default for blocks below 12 M entries; mutual patterns on the window kernel,
    // else the flags stay) | "flags" (default for larger blocks) |
    // "packets-fused" | "pull" | "flags" (A/B; the tests run all four)
    long long biggest = 0; // entries of the largest block
    for (int r = 0; r < P; ++r) biggest = std::max(biggest, (long long)rowptr[rs[(size_t)r + 1]] - (long long)rowptr[rs[(size_t)r]]);
    int proto = biggest < kDistPdlMinEntries ? 1 : 0; // short products: packets; long ones hide the flag chain (dist.py: the same rule)
    if (const char *env = getenv("RSG_DIST_EXCHANGE")) proto = !std::strcmp(env, "flags") ? 0 : !std::strcmp(env, "packets") ? 1 : !std::strcmp(env, "packets-fused") ? 3 : 2;
    if (P > 1 && rc == RSG_OK && proto == 2) {
        // ---- pull exchange (rsg_dist_enable_pull): per rank, the entries it reads, grouped by owner
        M->pull_col.assign((size_t)P, nullptr); M->pull_idx.assign((size_t)P, nullptr);
        for (int r = 0; r < P && rc == RSG_OK; ++r) {
            cudaSetDevice(M->dev[(size_t)r]);
            std::vector<int> col, idx, lo((size_t)P, 0), hi((size_t)P, 0);
            for (int o = 0; o < P; ++o) {
                lo[(size_t)o] = (int)col.size();
                if (o != r) {
                    const std::vector<int> &so = send[(size_t)o];
                    for (int c : need[(size_t)r])
                        if (c >= rs[(size_t)o] && c < rs[(size_t)o + 1]) {
                            const int li = c - rs[(size_t)o];
                            const int pos = (int)(std::lower_bound(so.begin(), so.end(), li) - so.begin());
                            col.push_back(M->hbase[(size_t)r] + o * width + pos);
                            idx.push_back(li);
                        }
                }
                hi[(size_t)o] = (int)col.size();
            }
            const size_t np = col.size();
            if (cudaMalloc(&M->pull_col[(size_t)r], sizeof(int) * (np ? np : 1)) != cudaSuccess || cudaMalloc(&M->pull_idx[(size_t)r], sizeof(int) * (np ? np : 1)) != cudaSuccess ||
                (np && (cudaMemcpy(M->pull_col[(size_t)r], col.data(), sizeof(int) * np, cudaMemcpyHostToDevice) != cudaSuccess ||
                        cudaMemcpy(M->pull_idx[(size_t)r], idx.data(), sizeof(int) * np, cudaMemcpyHostToDevice) != cudaSuccess))) {
                (void)cudaGetLastError(); rc = RSG_ALLOC_FAILED; break;
            }
            std::vector<const void *> px((size_t)(2 * P), nullptr);
            std::vector<const int *> pr((size_t)P, nullptr);
            for (int q = 0; q < P; ++q) {
                if (q == r) continue;
                if (std::find(nb[(size_t)r].begin(), nb[(size_t)r].end(), q) != nb[(size_t)r].end()) pr[(size_t)q] = M->flags[(size_t)q] + 160;
                if (hi[(size_t)q] > lo[(size_t)q])
                    for (int p = 0; p < 2; ++p) px[(size_t)(p * P + q)] = static_cast<char *>(M->x[(size_t)q]) + M->xs * (size_t)p * stride;
            }
            rc = rsg_dist_enable_pull(M->d[(size_t)r], (int)np, M->pull_col[(size_t)r], M->pull_idx[(size_t)r], P, px.data(), pr.data(), lo.data(), hi.data(),
                                      M->flags[(size_t)r] + 160, M->flags[(size_t)r] + 161);
            if (rc == RSG_UNSUPPORTED) { rc = RSG_OK; proto = 0; break; } // (a block on a gather kernel: everybody keeps the flags)
        }
        if (proto == 2) M->proto = 2;
        else for (int r = 0; r < P; ++r) if (M->d[(size_t)r]) M->d[(size_t)r]->pull = false;
    }
    bool want = P > 1 && rc == RSG_OK && (proto == 1 || proto == 3);
    for (int q = 0; q < P && want; ++q)
        for (int o : reads[(size_t)q])
            if (std::find(reads[(size_t)o].begin(), reads[(size_t)o].end(), q) == reads[(size_t)o].end()) want = false;
    for (int r = 0; r < P && want; ++r) {
        if (reads[(size_t)r].empty()) { want = false; break; }
        cudaSetDevice(M->dev[(size_t)r]);
        SpmvPlan *pl = nullptr;
        if (dist_plan(M->d[(size_t)r], &pl) != RSG_OK || !pl || pl->algo != RSG_SPMV_WINDOW) want = false;
    }
    if (want) {
        const size_t pbytes = 2 * M->xs, per_image = (size_t)P * (size_t)width;
        M->pk.assign((size_t)P, nullptr);
        for (int r = 0; r < P && rc == RSG_OK; ++r) {
            cudaSetDevice(M->dev[(size_t)r]);
            if (cudaMalloc(&M->pk[(size_t)r], 2 * per_image * pbytes) != cudaSuccess || cudaMemset(M->pk[(size_t)r], 0, 2 * per_image * pbytes) != cudaSuccess) {
                (void)cudaGetLastError(); rc = RSG_ALLOC_FAILED;
            }
        }
        for (int r = 0; r < P && rc == RSG_OK; ++r) { cudaSetDevice(M->dev[(size_t)r]); cudaDeviceSynchronize(); } // the zero fill precedes every mark
        for (int r = 0; r < P && rc == RSG_OK; ++r) {
            cudaSetDevice(M->dev[(size_t)r]);
            std::vector<void *> peer((size_t)(2 * P), nullptr);
            for (int q = 0; q < P; ++q)
                for (int p = 0; p < 2 && q != r; ++p)
                    peer[(size_t)(p * P + q)] = static_cast<char *>(M->pk[(size_t)q]) + pbytes * ((size_t)p * per_image + (size_t)r * (size_t)width);
            void *mine[2] = {M->pk[(size_t)r], static_cast<char *>(M->pk[(size_t)r]) + pbytes * per_image};
            rc = rsg_dist_enable_packets(M->d[(size_t)r], mine, peer.data(), proto == 1 ? 1 : 0);
        }
        for (int r = 0; r < P && rc == RSG_OK; ++r) { cudaSetDevice(M->dev[(size_t)r]); cudaDeviceSynchronize(); }
        M->packets = rc == RSG_OK;
        if (M->packets) M->proto = 1;
    }
    cudaSetDevice(prev);
    if (rc != RSG_OK) { multi_free(M); return rc; }
    *out = M;
    return RSG_OK;
}

int rsg_spmv_multi_destroy(rsg_multi_t M)
{
    if (!M) return RSG_BAD_ARG;
    int prev = 0;
    cudaGetDevice(&prev);
    multi_free(M);
    cudaSetDevice(prev);
    return RSG_OK;
}

/* x_host: n entries (fp64 for 'd' / 'm', fp32 for 's'); every rank's slice goes into both of its images */
int rsg_spmv_multi_set_x(rsg_multi_t M, const void *x_host)
{
    if (!M || (M->n > 0 && !x_host)) return RSG_BAD_ARG;
    const int rc = for_each_rank(M, [&](int r) -> int {
        const size_t nl = (size_t)M->n_local[(size_t)r];
        const char *src = static_cast<const char *>(x_host) + M->xs * (size_t)M->row_start[(size_t)r];
        for (int p = 0; p < 2 && nl; ++p)
            RSG_CUDA_TRY(cudaMemcpyAsync(static_cast<char *>(M->x[(size_t)r]) + M->xs * (size_t)p * M->stride, src, M->xs * nl,
                                         cudaMemcpyHostToDevice, M->stream[(size_t)r]));
        RSG_CUDA_TRY(cudaStreamSynchronize(M->stream[(size_t)r]));
        return RSG_OK;
    });
    return rc;
}

/* nsteps products y = A x on every device (asynchronous: returns once every device's launches are queued) */
int rsg_spmv_multi_step(rsg_multi_t M, int nsteps)
{
    if (!M || nsteps < 0) return RSG_BAD_ARG;
    return for_each_rank(M, [&](int r) -> int {
        for (int s = 0; s < nsteps; ++s) RSG_TRY(rsg_dist_step(M->d[(size_t)r], M->y[(size_t)r]));
        return RSG_OK;
    });
}

/* nsteps products timed on the devices (CUDA events on every device's stream, all devices released together):
 * *us_per_step = the slowest device's elapsed time / nsteps */
int rsg_spmv_multi_time(rsg_multi_t M, int nsteps, double *us_per_step)
{
    if (!M || nsteps < 1 || !us_per_step) return RSG_BAD_ARG;
    std::vector<float> ms((size_t)M->P, 0.f);
    RSG_TRY(for_each_rank(M, [&](int r) -> int { RSG_CUDA_TRY(cudaStreamSynchronize(M->stream[(size_t)r])); return RSG_OK; }));
    RSG_TRY(for_each_rank(M, [&](int r) -> int {
        RSG_CUDA_TRY(cudaEventRecord(M->ev0[(size_t)r], M->stream[(size_t)r]));
        for (int s = 0; s < nsteps; ++s) RSG_TRY(rsg_dist_step(M->d[(size_t)r], M->y[(size_t)r]));
        RSG_CUDA_TRY(cudaEventRecord(M->ev1[(size_t)r], M->stream[(size_t)r]));
        RSG_CUDA_TRY(cudaStreamSynchronize(M->stream[(size_t)r]));
        RSG_CUDA_TRY(cudaEventElapsedTime(&ms[(size_t)r], M->ev0[(size_t)r], M->ev1[(size_t)r]));
        return check_err_flag(M->h[(size_t)r]);
    }));
    float worst = 0.f;
    for (float v : ms) worst = std::max(worst, v);
    *us_per_step = (double)worst * 1e3 / nsteps;
    return RSG_OK;
}

/* waits for every device, then gathers y (m entries) to the host; RSG_DEPENDENCY_TIMEOUT if a halo wait gave up */
int rsg_spmv_multi_get_y(rsg_multi_t M, void *y_host)
{
    if (!M || (M->m > 0 && !y_host)) return RSG_BAD_ARG;
    return for_each_rank(M, [&](int r) -> int {
        const size_t nl = (size_t)M->n_local[(size_t)r];
        if (nl) RSG_CUDA_TRY(cudaMemcpyAsync(static_cast<char *>(y_host) + M->xs * (size_t)M->row_start[(size_t)r], M->y[(size_t)r], M->xs * nl,
                                             cudaMemcpyDeviceToHost, M->stream[(size_t)r]));
        return check_err_flag(M->h[(size_t)r]);
    });
}

/* Chained products x <- alpha * A * x, nsteps times, starting from x_host (n entries): x_host == NULL continues the chain.
 * us_per_step (may be NULL): device time of the slowest rank / nsteps.  RSG_UNSUPPORTED when some block has no chained form. */
int rsg_spmv_multi_chain(rsg_multi_t M, const void *x_host, int nsteps, double alpha, double *us_per_step)
{
    if (!M || nsteps < 0) return RSG_BAD_ARG;
    for (int r = 0; r < M->P; ++r) if (!M->dc[(size_t)r]) return RSG_UNSUPPORTED;
    if (x_host) { // image 1 of every rank, local entries AND halo, from the global vector; flags and step counters back to zero
        RSG_TRY(for_each_rank(M, [&](int r) -> int {
            cudaStream_t st = M->stream[(size_t)r];
            RSG_CUDA_TRY(cudaStreamSynchronize(st));
            const size_t nl = (size_t)M->n_local[(size_t)r];
            char *img1 = static_cast<char *>(M->x[(size_t)r]) + M->xs * M->stride;
            const char *src = static_cast<const char *>(x_host);
            RSG_CUDA_TRY(cudaMemcpyAsync(img1, src + M->xs * (size_t)M->row_start[(size_t)r], M->xs * nl, cudaMemcpyHostToDevice, st));
            const std::vector<int> &hs = M->halo_src[(size_t)r];
            std::vector<char> halo(hs.size() * M->xs, 0);
            for (size_t j = 0; j < hs.size(); ++j)
                if (hs[j] >= 0) std::memcpy(halo.data() + j * M->xs, src + M->xs * (size_t)hs[j], M->xs);
            if (!hs.empty()) RSG_CUDA_TRY(cudaMemcpyAsync(img1 + M->xs * (size_t)M->hbase[(size_t)r], halo.data(), halo.size(), cudaMemcpyHostToDevice, st));
            RSG_CUDA_TRY(cudaMemsetAsync(M->flags[(size_t)r] + 64, 0, sizeof(int) * 96, st));
            RSG_CUDA_TRY(cudaStreamSynchronize(st)); // (halo is a host temporary)
            return rsg_dist_chain_reset(M->dc[(size_t)r]);
        }));
    }
    std::vector<float> ms((size_t)M->P, 0.f);
    RSG_TRY(for_each_rank(M, [&](int r) -> int {
        cudaStream_t st = M->stream[(size_t)r];
        if (us_per_step) RSG_CUDA_TRY(cudaEventRecord(M->ev0[(size_t)r], st));
        for (int s = 0; s < nsteps; ++s) RSG_TRY(rsg_dist_step_chain(M->dc[(size_t)r], alpha));
        if (us_per_step) {
            RSG_CUDA_TRY(cudaEventRecord(M->ev1[(size_t)r], st));
            RSG_CUDA_TRY(cudaStreamSynchronize(st));
            RSG_CUDA_TRY(cudaEventElapsedTime(&ms[(size_t)r], M->ev0[(size_t)r], M->ev1[(size_t)r]));
            return check_err_flag(M->h[(size_t)r]);
        }
        return RSG_OK;
    }));
    if (us_per_step) {
        float worst = 0.f;
        for (float v : ms) worst = std::max(worst, v);
        *us_per_step = nsteps > 0 ? (double)worst * 1e3 / nsteps : 0.0;
    }
    return RSG_OK;
}

/* waits for every device, then gathers the current x of the chain (n entries) to the host */
int rsg_spmv_multi_get_x(rsg_multi_t M, void *x_host)
{
    if (!M || (M->n > 0 && !x_host)) return RSG_BAD_ARG;
    for (int r = 0; r < M->P; ++r) if (!M->dc[(size_t)r]) return RSG_UNSUPPORTED;
    return for_each_rank(M, [&](int r) -> int {
        const size_t nl = (size_t)M->n_local[(size_t)r];
        if (nl) RSG_CUDA_TRY(cudaMemcpyAsync(static_cast<char *>(x_host) + M->xs * (size_t)M->row_start[(size_t)r], rsg_dist_x(M->dc[(size_t)r]), M->xs * nl,
                                             cudaMemcpyDeviceToHost, M->stream[(size_t)r]));
        return check_err_flag(M->h[(size_t)r]);
    });
}

/* what: 0 = ranks, 1 = halo bytes all ranks together receive per step, 2 = SpMV kernel (rsg_spmv_algo) of rank 0,
 * 3 = halo width (entries per rank slot), 4 = exchange protocol (0 arrival flags, 1 packets, 2 pull); row_starts (may be NULL): ngpu+1 ints */
int rsg_spmv_multi_info(rsg_multi_t M, int what, long long *value, int *row_starts)
{
    if (!M || !value) return RSG_BAD_ARG;
    switch (what) {
    case 0: *value = M->P; break;
    case 1: *value = M->bytes_received_per_step; break;
    case 2: *value = M->h[0]->last_spmv_algo; break;
    case 3: *value = M->width; break;
    case 4: *value = M->proto; break;
    default: return RSG_BAD_ARG;
    }
    if (row_starts) for (int g = 0; g <= M->P; ++g) row_starts[g] = M->row_start[(size_t)g];
    return RSG_OK;
}

} // extern "C"
