// rsg_internal.cuh -- shared declarations of libreaspasol_gpu (sm_100a only).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <algorithm>
#include <atomic>
#include <vector>

#include "../../include/reaspasol_gpu.h"

#define RSG_CUDA_TRY(expr)                                         \
    do {                                                           \
        cudaError_t e__ = (expr);                                  \
        if (e__ != cudaSuccess) { (void)cudaGetLastError(); return RSG_CUDA_ERROR; } \
    } while (0)

#define RSG_TRY(expr)                      \
    do {                                   \
        int s__ = (expr);                  \
        if (s__ != RSG_OK) return s__;     \
    } while (0)

namespace rsg {

// cudaFuncSetAttribute applies to the CURRENT device only: one bit per device remembers what has been configured
// (a single process may drive several GPUs through several handles -- rsg_spmv_multi_*).
struct PerDeviceOnce {
    std::atomic<unsigned long long> mask{0};
    bool need(int device)
    {
        const unsigned long long bit = 1ull << (device & 63);
        return (mask.fetch_or(bit) & bit) == 0;
    }
};

// x-halo push folded into the SpMV launch (rsg_spmv_halo_fuse): the first `ctas` CTAs of the grid store
// x[send_idx[j]] into the peers' halo slots and release `epoch` to their flags, the rest are the tiles.
struct HaloFuse {
    void *dst[RSG_MAX_PEERS];
    int *flag[RSG_MAX_PEERS];
    int lo[RSG_MAX_PEERS], hi[RSG_MAX_PEERS]; // peer q reads only the entries j in [lo[q], hi[q]) of the send list
    const int *send_idx;
    int *counter;
    int nsend, npeers, epoch, ctas;
    // the send list as at most 4 contiguous runs of x (a slab of a stencil: its first and last plane): no index load
    int nruns, run_start[4], run_len[4];
    // packets != 0: dst[q] point at the peers' PACKET buffers and every entry travels as {value, epoch} in one store -- no fence,
    // no counter, no flag (rsg_dist_enable_packets); the readers poll the epoch inside the data
    int packets;
    // pull != 0 (rsg_dist_enable_pull): these CTAs FETCH instead -- dst[q] = peer q's x image (read), flag[q] = peer q's "x of
    // step e is final" counter (read, remote), lo/hi = owner q's range of the pull list, send_idx = index into the owner's x,
    // pull_col = where the entry goes in this rank's own image.  The product then contains no remote write at all.
    int pull;
    const int *pull_col;
    void *local_x;    // this rank's x image of the step (the halo part is written)
    int *my_ready;    // this rank's own counter, raised to `epoch` once the launch dependency is resolved
    int *local_flag;  // raised to `epoch` when the whole halo has been fetched: what the halo tiles wait for
};

// Receiving side of the packet exchange: pkts[c - halo_start] holds x entry c of this rank's halo as {value, epoch}
// (16 bytes for fp64 x: {lo, epoch, hi, epoch}; 8 bytes for fp32 x: {bits, epoch}).  pkts == nullptr: off.
struct PacketRecv {
    const void *pkts;
    int halo_start;
};

// Chained products x_{k+1} = alpha * A * x_k ("the SpMV epilogue stores its y rows straight into the peers' x
// buffers"): tiles that own rows a neighbour reads store those rows of y into the neighbours' halo slots of the NEXT x
// image as soon as they are computed, and the last of them to finish releases the step number to the neighbours' flags.
struct ChainPush {
    void *dst[RSG_MAX_PEERS];
    int *flag[RSG_MAX_PEERS];
    int lo[RSG_MAX_PEERS], hi[RSG_MAX_PEERS]; // peer q reads positions [lo[q], hi[q]) of the send list
    const int *send_pos;                      // per local row: its position in the send list, or -1
    int *counter;                             // producer tiles finished (reset by the last one)
    int npeers, nprod, epoch;
};

// Long rows of a TILE plan (longer than a CTA's staging capacity) are cut into segments, one CTA each, placed in
// the grid ahead of the tiles: partial sums meet in `part`, the last segment of a row to arrive adds them in
// segment order (deterministic) and stores y.  The tiles skip those rows.
struct LongSplit {
    const int4 *seg;  // per segment: (row, first zero-based entry, end entry, index into lrow)
    const int2 *lrow; // per long row: (first segment slot, number of segments)
    double *part;     // one 8-byte slot per segment
    int *cnt;         // arrivals per long row (reset by the last arrival)
    int nseg;
    unsigned long long *trace; // debug (rsg_debug_trace): 4 x %globaltimer per CTA, as in the window kernel (bit 63 of word 3: segment CTA)
};

// ------------------------------------------------------------------ SpMV plan
// Built once per sparsity pattern by rsg_spmv_analysis (device kernels + one small D2H of
// the statistics), cached in the handle; the analogue of cusparseSpMV_bufferSize/preprocess.
struct SpmvPlan {
    // key
    const int *rowptr = nullptr;
    const int *colidx = nullptr;
    int m = 0, n = 0, base = 0;
    int nnz_hint = 0;      // the caller's nnz argument (part of the key: pointers get recycled)
    int opt_sig = 0;       // hash of the options that shape the plan
    // statistics
    int nnz = 0;           // stored entries, rowptr[m]-rowptr[0]
    int first = 0;         // rowptr[0]-base (non-zero for a row-slab view into a larger CSR)
    bool scattered = false; // neighbouring rows read unrelated x lines (scatter statistic > 0.5): the tile kernel stages products
    int max_row = 0;
    int algo = 0;          // resolved rsg_spmv_algo
    int variant = 0;       // resolved tile/CTA shape
    // row-block kernel
    int tile = 0;          // merged items (rows + entries) per CTA
    int nblk = 0;
    int *blk_row = nullptr;    // nblk+1 row indices
    int4 *pdesc = nullptr;     // tile kernel: nblk+1 (first row, zero-based entry offset, longest row, halo mark)
    int *rp0 = nullptr;        // tile / window kernels: zero-based padded copy of rowptr
    int rp0_len = 0;
    // tile kernel: split long rows (LongSplit)
    int4 *seg = nullptr;
    int2 *lrow = nullptr;
    double *seg_part = nullptr;
    int *seg_cnt = nullptr;
    int nseg = 0, nlong = 0;
    // window kernel (spmv_window.cu): 128-byte tile records, 16-bit window-local column indices
    int4 *wdesc = nullptr;
    uint16_t *widx = nullptr;
    int wsmem = 0, wthreads = 0, whalo = 0, first8 = 0;
    int wprod = 0;             // window plan built for chained products: tiles that own rows a neighbour reads
    int vbytes = 0, xbytes = 0; // stored-value / vector element sizes the plan was shaped for
    int lanes_hint = 0;        // tile kernel: 2 = the stage holds about half as many rows as the CTA has threads -> two lanes per row
    // merge-path kernel
    int merge_items = 0;   // rows+entries per CTA
    int ntiles = 0;
    int2 *coords = nullptr;    // ntiles+1 (row, entry) start coordinates
    int *carry_row = nullptr;  // ntiles
    void *carry_val = nullptr; // ntiles * 8 B
    size_t bytes = 0;
};

// ------------------------------------------------ triangular / ILU0 plan
struct TriPlan {
    const int *rowptr = nullptr;
    const int *colidx = nullptr;
    int n = 0, base = 0, nnz = 0;
    int nnz_hint = 0;              // the caller's nnz argument (part of the key)
    int *diag_pos = nullptr;       // zero-based position of a_ii in colidx/values, -1 if missing
    int missing_diag_row = -1;     // first row without a diagonal entry
    bool unsorted = false;         // some row's columns are not ascending: ILU0 / SpTRSV refuse it (RSG_BAD_ARG)
    int max_row = 0;
    bool have[2] = {false, false}; // [0]=L (lower pattern), [1]=U (upper pattern)
    int nlevels[2] = {0, 0};
    int *lev[2] = {nullptr, nullptr};       // level of every row
    int *order[2] = {nullptr, nullptr};     // rows sorted by (level, row): "level_rows"
    int *level_ptr[2] = {nullptr, nullptr}; // nlevels+1 offsets into order (device)
    int *crit[2] = {nullptr, nullptr};      // per row: the dependency (zero-based column) with the deepest level, -1 if none
    std::vector<int> level_ptr_host[2];
    int *done = nullptr;           // per-row dependency flags (epoch-stamped)
    int epoch = 0;
    int *ctl = nullptr;            // 16 ints: [0] ticket, [1] error flag, [2] min zero-pivot row; [4..7] second kernel
    int *sort_keys = nullptr;      // radix-sort scratch (n ints each)
    int *sort_vals = nullptr;
    // chain solves (ilu0_trsv.cu: trsv_chain_kernel): per position of the level order (row, first entry, end entry, diagonal
    // position), and the (first position | new-level flag << 31, end position) chunks of at most kChainThreads rows
    int4 *tmeta[2] = {nullptr, nullptr};
    int2 *chunks[2] = {nullptr, nullptr};
    int nchunks[2] = {0, 0};
    int *slab = nullptr;           // ONE allocation holding every array above (re-used when the analysis is re-run)
    int slab_n = 0;
    size_t bytes = 0;
};

// ------------------------------------------------ transposed pattern (CSC) of a CSR matrix, cached per handle
struct TransposePlan {
    const int *rowptr = nullptr;
    const int *colidx = nullptr;
    int m = 0, n = 0, base = 0, nnz_hint = 0;
    int nnz = 0, first = 0;
    int *colptr = nullptr;    // n+1, carries the base
    int *rowidx = nullptr;    // nnz, carries the base
    int *perm = nullptr;      // nnz: entry (relative to `first`) that lands at k
    void *values_t = nullptr; // nnz * 8 bytes, re-gathered per product
    const void *gathered_from = nullptr; // values pointer (and element size) values_t currently mirrors:
    int gathered_size = 0;               // consulted only under RSG_OPT_TRANSPOSE_STATIC
    int max_col = -1;                    // longest column of A (-1 unknown)
};

} // namespace rsg

struct rsg_context {
    int device = 0;
    cudaStream_t stream = nullptr;
    int sm_count = 148;
    int l2_bytes = 0;
    int max_smem_optin = 0;
    int opts[RSG_OPT_COUNT_] = {0};
    long long launches = 0;
    int last_spmv_algo = 0;
    int last_spmv_blocks = 0;
    int last_spmv_maxrow = 0;
    std::vector<rsg::SpmvPlan> spmv_plans;
    std::vector<rsg::TriPlan> tri_plans;
    std::vector<rsg::TransposePlan> transpose_plans;
    int last_tri = -1;
    void *cub_tmp = nullptr;
    size_t cub_tmp_bytes = 0;
    int *scratch_i = nullptr; // small device scratch (stats), 64 ints
    int *err_flag = nullptr;  // = scratch_i + 60: raised by a device-side dependency wait that gave up
    int *pinned_i = nullptr;  // small pinned host mirror, 64 ints
    const int *wait_flags = nullptr;   // one-shot halo wait for the next SpMV (rsg_spmv_halo_wait)
    int wait_n = 0, wait_epoch = 0;
    rsg::ChainPush chain = {};         // one-shot epilogue push for the next SpMV (rsg_dist_step_chain); npeers == 0: none
    const int *plan_send_pos = nullptr; // read when a window plan is built: marks producer tiles, orders boundary tiles first
    unsigned long long *trace = nullptr; // rsg_debug_trace: per-CTA timestamps of the window kernel (debug; nullptr = off)
    int trace_ctas = 0, trace_parity = 0;
    rsg::PacketRecv packets = {nullptr, 0}; // one-shot: the next SpMV's halo tiles read their halo entries out of packets
    rsg::HaloFuse fuse = {};           // one-shot fused push for the next SpMV (rsg_spmv_halo_fuse); ctas == 0: none
    const void *persist_ptr = nullptr; // x currently covered by the L2 access-policy window
    size_t persist_bytes = 0;
    void *krylov_ws = nullptr;         // grow-only workspace of rsg_pcg_d (vectors, partial sums, scalars)
    size_t krylov_bytes = 0;
    double *krylov_pin = nullptr;      // pinned host mirror of the scalars
};

struct rsg_matrix {
    char precision;
    int m, n, nnz, base;
    int *rowptr;
    int *colidx;
    void *values;
    void *x;
    void *y;
    void *pin_x; // pinned staging
    void *pin_y;
    // pipelined host-vector product (rsg_matrix_spmv_host): row chunks fixed at upload time
    static constexpr int kMaxChunks = 32;
    int nchunk;                   // 1 = not pipelined
    int chunk_row[kMaxChunks + 1]; // first row of every chunk
    int chunk_x[kMaxChunks + 1];   // chunk c may start once x[0 .. chunk_x[c+1]) is on the device (monotone)
    cudaStream_t s_in, s_out, s_main; // x upward, y downward, products (forked from / joined into s_main)
    cudaEvent_t ev_start, ev_in[kMaxChunks], ev_mul[kMaxChunks], ev_pre, ev_join_in, ev_join_out;
    // the whole pipelined step as one CUDA graph, captured on the second call with the same buffers and scalars
    cudaGraphExec_t graph;
    const void *g_x, *g_y, *seen_x, *seen_y;
    double g_alpha, seen_alpha;
    int g_sig, seen_sig, g_kernels;
};

namespace rsg {

int ensure_cub_tmp(rsg_context *h, size_t bytes);

// spmv.cu
// pkey: stored-value bytes | vector-element bytes << 4 (0x88 fp64, 0x44 fp32, 0x84 mixed; a bare 8 / 4 means both alike)
int spmv_get_plan(rsg_context *h, int m, int n, int nnz_hint, int base, const int *rowptr, const int *colidx,
                  bool force_rebuild, SpmvPlan **out, int pkey = 8);
// polls a device-side flag wait may spend before it gives up (RSG_OPT_WAIT_SPINS; default 2^22: seconds)
inline int wait_spins(const rsg_context *h) { const int v = h->opts[RSG_OPT_WAIT_SPINS]; return v > 0 ? v : (1 << 22); }
// RSG_DEPENDENCY_TIMEOUT if a kernel of this handle raised the flag since the last check (synchronises the stream)
int check_err_flag(rsg_context *h);
void spmv_free_plan(SpmvPlan &p);
long long spmv_plans_freed();
template <typename TV, typename TX>
int spmv_run(rsg_context *h, int m, int n, int nnz_hint, int base, const int *rowptr, const int *colidx,
             const TV *values, TX alpha, const TX *x, TX beta, TX *y);
int row_partition(rsg_context *h, int m, int base, const int *rowptr, int ngpu, int balanced,
                  int *row_starts_host);

// spmv_tile.cu
int tile_num_variants();
template <typename TV, typename TX>
int tile_launch(rsg_context *h, const SpmvPlan *p, const TV *values, TX alpha, const TX *x, TX beta, TX *y);
int pipe_build_plan(rsg_context *h, SpmvPlan &p, bool strict_split = false);

// spmv_window.cu
int window_build_plan(rsg_context *h, SpmvPlan &p, int vbytes, int xbytes, bool strict);
int window_place_halo(rsg_context *h, SpmvPlan &p, int start_pct, int end_pct); // where the halo tiles sit in the grid (see spmv_window.cu)
template <typename TV, typename TX>
int window_launch(rsg_context *h, const SpmvPlan *p, const TV *values, TX alpha, const TX *x, TX beta, TX *y);

// transpose.cu
void transpose_free_plan(TransposePlan &p);
void transpose_drop_plan(rsg_context *h, size_t i);
template <typename T>
int csr2csc_run(rsg_context *h, int m, int n, int base, const int *rowptr, const int *colidx, const T *values, int *colptr,
                int *rowidx, T *values_out);
template <typename T>
int spmv_transposed_run(rsg_context *h, int m, int n, int nnz_hint, int base, const int *rowptr, const int *colidx, const T *values,
                        T alpha, const T *x, T beta, T *y);

// levels.cu
int tri_get_plan(rsg_context *h, int n, int nnz_hint, int base, const int *rowptr, const int *colidx,
                 bool force_rebuild, TriPlan **out);
int tri_build_levels(rsg_context *h, TriPlan *p, int which); // which = -1: both, one launch
void tri_free_plan(TriPlan &p);

// ilu0_trsv.cu
template <typename T>
int ilu0_run(rsg_context *h, TriPlan *p, T *values, int *zero_pivot_row);
template <typename T>
int trsv_run(rsg_context *h, TriPlan *p, const T *lu, const T *b, T *z, T *y);

} // namespace rsg
