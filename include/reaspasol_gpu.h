/*
 * reaspasol_gpu.h -- C ABI of libreaspasol_gpu: B200-native CSR SpMV, ILU0 and L/U SpTRSV.
 *
 * The reference (yex33/ReSpaSol) has no plugin layer: its hot path is the set of cuSPARSE
 * calls inside two main()s.  Each entry point below replaces one of those call groups and
 * takes the same raw device CSR arrays (int32 indices, base 0 or 1) the reference harness
 * uploads.  Citations are relative to /root/reference.
 *
 *   spmv_csr_d / spmv_csr_s      cusparseSpMV fp64 / fp32           GPU/spmv.cu:184-186 / 179-181
 *                                (+ descriptors & workspace)         GPU/spmv.cu:122-164
 *   spmv_csr_mixed               (no reference equivalent; north_star: fp32 A, fp64 x/y/accumulate)
 *   rsg_ilu0_analysis            cusparse?csrilu02_analysis + zeroPivot   GPU/ilu0.cu:196-226
 *   ilu0_factor_d / _s           cusparse?csrilu02 + zeroPivot            GPU/ilu0.cu:257-282
 *   rsg_sptrsv_analysis          cusparse?csrsv2_analysis x2              GPU/ilu0.cu:229-252
 *   sptrsv_lu_d / _s             cusparse?csrsv2_solve x2                 GPU/ilu0.cu:284-310
 *                                (second solve = the INTENDED U solve; the reference's call is an
 *                                 L^T solve left over from an IC0 example -- SURVEY.md 9-D3)
 *   *_host wrappers              cudaMalloc + cudaMemcpy H2D/D2H          GPU/spmv.cu:91-119,216
 *                                                                         GPU/ilu0.cu:78-118
 *
 * Conventions
 *   - All pointers named rowptr/colidx/values/x/y/b/z are DEVICE pointers unless the function
 *     name ends in _host.  Calls are asynchronous on the handle's stream (rsg_set_stream);
 *     analysis functions and anything returning a value through a host pointer synchronise.
 *   - The caller owns every array; the library owns only plans/workspace inside the handle
 *     (the analogue of dBuffer / pBuffer, spmv.cu:164, ilu0.cu:193-194).
 *   - `nnz` is advisory: the stored-entry count is rowptr[m]-rowptr[0].  When they disagree
 *     (the reference loader's symmetric-file defect) the rowptr value is used.
 *   - ILU0 and SpTRSV need the columns of every row in ascending order (what the loader produces); an unsorted row makes
 *     the analysis, factor and solve calls return RSG_BAD_ARG.  SpMV accepts any order.
 *   - One handle = one device + one stream; calls on a handle are not re-entrant.
 *   - Every function returns an rsg_status (0 = OK).  The library never exit()s or prints.
 */
#ifndef REASPASOL_GPU_H
#define REASPASOL_GPU_H

#ifdef __cplusplus
extern "C" {
#endif

typedef struct rsg_context *rsg_handle_t;
typedef struct rsg_matrix *rsg_matrix_t;
typedef struct rsg_dist *rsg_dist_t;   /* one rank's side of the row-block product (one process or thread per GPU) */
typedef struct rsg_multi *rsg_multi_t; /* a whole matrix over N devices of one process */

typedef enum {
    RSG_OK = 0,
    RSG_BAD_ARG = 1,
    RSG_CUDA_ERROR = 2,
    RSG_ALLOC_FAILED = 3,
    RSG_NOT_ANALYSED = 4,
    RSG_ZERO_PIVOT_STRUCTURAL = 5, /* A(j,j) missing  -- ilu0.cu:221-226 */
    RSG_ZERO_PIVOT_NUMERIC = 6,    /* U(j,j) == 0     -- ilu0.cu:278-282 */
    RSG_DEPENDENCY_TIMEOUT = 7,    /* watchdog in a dependency-flag wait fired (malformed input) */
    RSG_UNSUPPORTED = 8,
    RSG_BREAKDOWN = 9,             /* Krylov recurrence broke down (a scalar became 0, inf or NaN) */
    RSG_SIZE_LIMIT = 10            /* the matrix is larger than the fixed scratch of an allocation-free entry point */
} rsg_status;

/* ---- handle ---------------------------------------------------------------------------- */
int rsg_create(rsg_handle_t *handle, int device);
int rsg_destroy(rsg_handle_t handle);
int rsg_set_stream(rsg_handle_t handle, void *cuda_stream); /* cudaStream_t; NULL = default */
int rsg_synchronize(rsg_handle_t handle); /* RSG_DEPENDENCY_TIMEOUT if a device-side wait gave up since the last call */
const char *rsg_status_string(int status);
int rsg_version(void);

/* ---- tuning / introspection ------------------------------------------------------------- */
typedef enum {
    RSG_OPT_SPMV_ALGO = 0,      /* rsg_spmv_algo; default AUTO */
    RSG_OPT_SPMV_REDUCE = 1,    /* row-block kernel's in-tile reduction: 0 auto, 1 thread/row, 2 warp/row */
    RSG_OPT_SPMV_LONG_ROW = 2,  /* rows longer than this get a CTA each (0 = default, tile/4) */
    RSG_OPT_SPMV_VARIANT = 3,   /* compiled tile/CTA shape variant index (0 = default) */
    RSG_OPT_SPMV_CACHE_HINT = 4,/* 0: plain ld.global.nc; 1: L1::no_allocate streams; 2: + L2 evict_first streams */
    RSG_OPT_TRSV_ALGO = 5,      /* 0 auto (by level width), 1 sync-free flags on every SM, 2 level-scheduled launches,
                                   3 chain: one CTA walks the levels, a block barrier per level (narrow levels) */
    RSG_OPT_ILU0_ALGO = 6,      /* 0 auto, 1 sync-free flags, 2 level-scheduled launches */
    RSG_OPT_SPMV_X_PERSIST = 7, /* 1: pin x in L2 with an access-policy window on the stream */
    RSG_OPT_SPMV_X_HINT = 8,    /* A/B and timing switches (0 = default): 4 = tile plan never sized for two lanes per row; 2 = tile kernel
                                   without product staging on scattered patterns; bits 8 / 16 / 32 / 64 / 128 = rsg_dist_step timing
                                   experiments of tools/dist_trace.py (no halo wait / no push / never plain launches / fewer or more
                                   push CTAs) -- bits 8 and 16 give WRONG results, never set them outside a timing run */
    RSG_OPT_SPMV_ROWS_PER_THREAD = 9, /* tile kernel: tile size in rows per thread (0 = 1) */
    RSG_OPT_SPMV_CTAS_PER_SM = 10,    /* unused since the persistent kernel was retired */
    RSG_OPT_TRSV_BATCH = 11,    /* sync-free solves: solution entries polled per round trip (1,2,4,8,16; 0 = default) */
    RSG_OPT_TRSV_GATE = 12,     /* sync-free solves: 1 = a row first polls only its deepest dependency, 2 = off, 0 = default */
    RSG_OPT_TRI_CTAS_PER_SM = 13, /* sync-free ILU0 / solves: cap on resident CTAs per SM (0 = default) */
    RSG_OPT_TRSV_BLOCK = 14,    /* sync-free solves: threads per CTA (multiple of 32, <= 256; 0 = default) */
    RSG_OPT_SPMV_HALO_START = 15, /* multi-GPU: zero-based column from which x is filled by peers (0 = none); read when
                                     the SpMV plan is built: tiles that touch such columns are marked and can wait on
                                     arrival flags (rsg_spmv_halo_wait) */
    RSG_OPT_HOST_CHUNKS = 16,   /* rsg_matrix_upload: row chunks of the pipelined host-vector product
                                   (x H2D | SpMV | y D2H overlap chunk by chunk); 0 = default (4), 1 = one piece */
    RSG_OPT_HOST_Y_DIRECT = 17, /* pipelined host-vector product: 1 = the kernels store y straight into the page-locked host
                                   buffer (mapped memory, no D2H copy stage); 0 = device memory + async D2H copies (default) */
    RSG_OPT_HOST_GRAPH = 18,    /* pipelined host-vector product: 0/1 = replay the step from a CUDA graph once the same host buffers
                                   and alpha are seen twice (one launch call instead of ~30 stream calls); 2 = never */
    RSG_OPT_TRANSPOSE_STATIC = 19, /* spmv_csr_t_*: 1 = the caller promises the values behind a given `values` pointer do not
                                   change between calls (an iterative solver on a fixed matrix): the transposed copy is
                                   gathered once per (pattern, values pointer, precision) and the product costs what the
                                   untransposed one does; 0 = read the values through the cached permutation on every call
                                   (default, always right: fused kernel, or gather + product when a column is longer
                                   than 2048 entries); 2 = always gather + product (A/B) */
    RSG_OPT_SPMV_WIN_BYTES = 20, /* window kernel: shared memory per CTA the plan aims at (bytes; 0 = default 36 KB) */
    RSG_OPT_SPMV_WIN_STREAM_PCT = 21, /* window kernel: share of that given to the entries (values + 16-bit columns), percent (0 = 64) */
    RSG_OPT_SPMV_WIN_GAP = 22,  /* window kernel plan: runs of x separated by up to this many unused 4-column granules are merged (0 = 3) */
    RSG_OPT_WAIT_SPINS = 23,    /* polls a device-side dependency / halo wait may spend before it gives up and the next
                                   synchronising call returns RSG_DEPENDENCY_TIMEOUT (0 = default 2^22, i.e. seconds) */
    RSG_OPT_SPMV_PDL = 24,      /* window kernel, programmatic dependent launch: 0 (default) = a product's CTAs may start while the
                                   previous kernel of the stream drains, reading only plan-owned data (tile records, row pointers,
                                   16-bit columns) before they wait for it; 2 = the matrix VALUES too -- the caller promises that
                                   nothing queued ahead of the product writes them (a solver loop on a fixed matrix, the harness's
                                   50 timed products, spmv.cu:174-207); 1 = plain launches */
    RSG_OPT_SPMV_HALO_PLACE = 25, /* multi-GPU window plans: share of the grid (percent) that is interior tiles before the first tile
                                   that reads a peer's entries; from there the two kinds alternate, the last wave is interior again.
                                   0 = default (packet exchange: 20, ending at 60; arrival flags: 100 = all such tiles at the end).  Read when the
                                   plan is built and by rsg_dist_enable_packets; not part of the plan's key */
    RSG_OPT_SPMV_HALO_END = 26,  /* the same plans: share of the grid (percent) by which the last such tile has been placed (0 = default:
                                   all but the last wave) */
    RSG_OPT_COUNT_
} rsg_option;

typedef enum {
    RSG_SPMV_AUTO = 0,
    RSG_SPMV_ROWBLOCK = 1, /* product-staged row blocks (register-path loads): thread-, warp-, CTA-per-row */
    RSG_SPMV_MERGE = 2,    /* merge-path, for skewed row lengths */
    RSG_SPMV_VECTOR = 3,   /* plain sub-warp-per-row, scalar loads (any alignment) */
    RSG_SPMV_SCALAR = 4,   /* plain thread-per-row, scalar loads */
    RSG_SPMV_STREAM = 5,   /* retired (round 1's first TMA kernel): selects RSG_SPMV_TILE */
    RSG_SPMV_PIPE = 6,     /* retired (persistent multi-stage TMA ring, measured slower everywhere): selects RSG_SPMV_TILE */
    RSG_SPMV_TILE = 7,     /* one TMA-staged tile per CTA (row pointers staged too), many small CTAs per SM; long rows split */
    RSG_SPMV_WINDOW = 8    /* TMA-staged tile + its x window in shared memory, 16-bit window-local columns built at plan time
                              (stencil / FEM / banded patterns; falls back to TILE when a pattern has no compact windows) */
} rsg_spmv_algo;

typedef enum {
    RSG_INFO_KERNEL_LAUNCHES = 0, /* kernels of this library launched on the handle so far */
    RSG_INFO_SPMV_ALGO_USED = 1,  /* rsg_spmv_algo of the last spmv call */
    RSG_INFO_SPMV_ROW_BLOCKS = 2, /* CTAs in the last spmv plan */
    RSG_INFO_SPMV_MAX_ROW = 3,    /* longest row seen by the last spmv analysis */
    RSG_INFO_LEVELS_L = 4,
    RSG_INFO_LEVELS_U = 5,
    RSG_INFO_SM_COUNT = 6,
    RSG_INFO_WORKSPACE_BYTES = 7
} rsg_info;

/* Plans (SpMV tiling, level schedules) are cached per handle, keyed on the device addresses of
 * rowptr/colidx plus m, n, nnz and base.  When a matrix is freed and another one may be allocated
 * at the same addresses, call rsg_forget_matrix (or re-run the analysis, which always rebuilds). */
int rsg_forget_matrix(rsg_handle_t handle, const int *rowptr);
/* Debug aid (tools/dist_trace.py): while buf != NULL the window kernel writes four %globaltimer values (ns) per CTA of every
 * product into the DEVICE array buf (2 * 4 * capacity_ctas words: two halves used in turn by consecutive products) at
 * [half][4 * blockIdx]: entry, launch dependency resolved, tile staged in shared memory, done (bit 63 of the last word: the
 * tile reads a peer's entries).  Products with more CTAs than capacity_ctas are not traced. */
int rsg_debug_trace(rsg_handle_t handle, unsigned long long *buf, int capacity_ctas);
int rsg_set_option(rsg_handle_t handle, int option, int value);
int rsg_get_info(rsg_handle_t handle, int info, long long *value);

/* ---- SpMV:  y = alpha*A*x + beta*y  (beta == 0 never reads y) ---------------------------- */
int rsg_spmv_analysis(rsg_handle_t handle, int m, int n, int nnz, int base,
                      const int *rowptr, const int *colidx);
int spmv_csr_d(rsg_handle_t handle, int m, int n, int nnz, int base,
               const int *rowptr, const int *colidx, const double *values,
               double alpha, const double *x, double beta, double *y);
int spmv_csr_s(rsg_handle_t handle, int m, int n, int nnz, int base,
               const int *rowptr, const int *colidx, const float *values,
               float alpha, const float *x, float beta, float *y);
int spmv_csr_mixed(rsg_handle_t handle, int m, int n, int nnz, int base,
                   const int *rowptr, const int *colidx, const float *values,
                   double alpha, const double *x, double beta, double *y);

/* Allocation-free fp64 SpMV: y = alpha*A*x + beta*y.  NO counterpart in the reference; added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * it).  Takes no handle, runs on the CURRENT device, on `cuda_stream` (cudaStream_t; NULL = default stream), and calls no
 * cudaMalloc / cudaFree / allocating routine: the tile kernel's plan (tile descriptors + padded row pointers) is rebuilt on
 * every call in a file-scope device array of RSG_NOALLOC_SCRATCH_BYTES (about 5 MB) that every GPU loading the library holds
 * until exit.  Limits: m <= RSG_NOALLOC_MAX_ROWS and at most RSG_NOALLOC_MAX_TILES tiles (tiles = ceil((4*m + nnz) / t),
 * 512 <= t <= 2048 chosen from nnz/m; 4*m + nnz <= 33 554 432 always fits) -- beyond that RSG_SIZE_LIMIT, nothing launched.
 * Contract: `nnz` is EXACT here (= rowptr[m] - rowptr[0]; nothing is read back to check it) and rowptr[0] == base; colidx and
 * values 16-byte aligned (RSG_BAD_ARG otherwise); asynchronous, four kernel launches.  All calls share the one array: calls
 * on one stream are ordered, calls that OVERLAP on a GPU (two streams, two host threads) race on it and give wrong results.
 * Results agree with spmv_csr_d within its tolerance (1e-12 * max|y|; the summation order inside a row may differ). */
#define RSG_NOALLOC_MAX_ROWS (1 << 20)
#define RSG_NOALLOC_MAX_TILES (1 << 16)
#define RSG_NOALLOC_SCRATCH_BYTES 5243392
int spmv_csr_d_noalloc(int m, int n, int nnz, int base, const int *rowptr, const int *colidx, const double *values,
                       double alpha, const double *x, double beta, double *y, void *cuda_stream);

/* ---- transpose / CSC path ------------------------------------------------- */
/* CSR -> CSC on the device: what the reference obtains by asking its loader for the transposed matrix
 * (transpose = 1, loadMatrixMarket.cpp:79-81,124; test_superLU_MT.c:85).  colptr (n+1 ints), rowidx and values_out
 * (stored-count entries each) are caller-allocated DEVICE arrays; indices keep `base`; inside a column the entries
 * keep their row order (stable).  values / values_out may be NULL (pattern only).  Synchronous. */
int rsg_csr2csc_d(rsg_handle_t handle, int m, int n, int nnz, int base, const int *rowptr, const int *colidx,
                  const double *values, int *colptr, int *rowidx, double *values_out);
int rsg_csr2csc_s(rsg_handle_t handle, int m, int n, int nnz, int base, const int *rowptr, const int *colidx,
                  const float *values, int *colptr, int *rowidx, float *values_out);
/* y = alpha*A^T*x + beta*y (x has m entries, y has n): CUSPARSE_OPERATION_TRANSPOSE of the SpMV call group
 * (spmv.cu:184 passes NON_TRANSPOSE; this is the CSC product).  The transposed pattern is built once per matrix and
 * cached in the handle; the values are read through that permutation on every call (unless RSG_OPT_TRANSPOSE_STATIC = 1,
 * see there): deterministic, no atomics. */
int spmv_csr_t_d(rsg_handle_t handle, int m, int n, int nnz, int base, const int *rowptr, const int *colidx,
                 const double *values, double alpha, const double *x, double beta, double *y);
int spmv_csr_t_s(rsg_handle_t handle, int m, int n, int nnz, int base, const int *rowptr, const int *colidx,
                 const float *values, float alpha, const float *x, float beta, float *y);

/* ---- ILU0 -------------------------------------------------------------------------------- */
/* "Symbolic": diagonal positions, structural zero-pivot check, level sets of the lower pattern
 * (shared by the factorization and the L solve).  *zero_pivot_row = -1 or the first row with no
 * diagonal entry (then the status is RSG_ZERO_PIVOT_STRUCTURAL). */
int rsg_ilu0_analysis(rsg_handle_t handle, int n, int nnz, int base,
                      const int *rowptr, const int *colidx, int *zero_pivot_row);
/* "Numeric": in-place ILU(0), natural order, no pivoting; runs the analysis itself if needed.
 * zero_pivot_row may be NULL (then the call stays asynchronous and pivots are not reported). */
int ilu0_factor_d(rsg_handle_t handle, int n, int nnz, int base,
                  const int *rowptr, const int *colidx, double *values_inout, int *zero_pivot_row);
int ilu0_factor_s(rsg_handle_t handle, int n, int nnz, int base,
                  const int *rowptr, const int *colidx, float *values_inout, int *zero_pivot_row);

/* ---- SpTRSV:  L z = b (unit lower), U y = z (non-unit upper) on the in-place LU values ---- */
int rsg_sptrsv_analysis(rsg_handle_t handle, int n, int nnz, int base,
                        const int *rowptr, const int *colidx);
int sptrsv_lu_d(rsg_handle_t handle, int n, int nnz, int base,
                const int *rowptr, const int *colidx, const double *lu_values,
                const double *b, double *tmp_z, double *y);
int sptrsv_lu_s(rsg_handle_t handle, int n, int nnz, int base,
                const int *rowptr, const int *colidx, const float *lu_values,
                const float *b, float *tmp_z, float *y);

/* ---- parity introspection (integers must be bit-exact against the oracle) ----------------- */
/* which: 0 = L (lower pattern), 1 = U (upper pattern).  HOST output buffers: level_ptr needs
 * nlevels+1 ints, level_rows n ints; pass NULL for both to query *nlevels only. */
int rsg_get_levels(rsg_handle_t handle, int which, int *nlevels, int *level_ptr, int *level_rows);
/* row_starts: HOST buffer of ngpu+1 ints.  balanced=1: nnz-balanced split points
 * start[g] = first row r with rowptr[r]-rowptr[0] >= floor(g*nnz/ngpu); 0: ceil(g*m/ngpu). */
int rsg_get_row_partition(rsg_handle_t handle, int m, int base, const int *rowptr,
                          int ngpu, int balanced, int *row_starts);

/* ---- multi-GPU x-halo push (north_star: row-block SpMV across the box; no reference equivalent) -- */
/* x_local[send_idx[j]] (j < nsend) is stored to peer_slots[q][j] for every q < npeers with a non-NULL
 * slot.  peer_slots is a HOST array of DEVICE pointers into the peers' halo buffers (P2P-mapped, e.g.
 * CUDA IPC / symmetric memory), each already offset to this rank's slot.  Asynchronous on the handle's
 * stream; the caller orders it against the peers' reads with its own cross-GPU barrier. */
#define RSG_MAX_PEERS 16
/* peer_flags (may be NULL): HOST array of npeers DEVICE pointers to one int each, in the peers' memory; after
 * every store of this launch is visible system-wide, `epoch` is stored (release, system scope) to each
 * non-NULL flag -- "this rank's entries for step `epoch` have landed". */
int rsg_halo_push(rsg_handle_t handle, int elem_bytes, int nsend, const int *send_idx, const void *x_local,
                  int npeers, void *const *peer_slots, int *const *peer_flags, int epoch);
/* One-shot: the NEXT spmv_csr_* call on this handle makes every tile that reads columns >=
 * RSG_OPT_SPMV_HALO_START wait (acquire, system scope) until flags[i] >= epoch for all i < nflags
 * (flags: DEVICE pointer, nflags <= RSG_MAX_PEERS), while all other tiles run at once: the exchange
 * overlaps the product without splitting it.  Needs the TILE kernel (AUTO picks it for regular rows);
 * (the last tile waits too, marked or not, so a rank never runs more than one step ahead of a neighbour).  Needs the
 * WINDOW or TILE kernel (AUTO picks them for regular rows); other kernels return RSG_UNSUPPORTED.  A wait that gives up
 * (RSG_OPT_WAIT_SPINS polls) raises a flag that rsg_synchronize returns as RSG_DEPENDENCY_TIMEOUT. */
int rsg_spmv_halo_wait(rsg_handle_t handle, const int *flags, int nflags, int epoch);
/* One-shot: the NEXT spmv_csr_* call also performs the push of rsg_halo_push itself -- its first few CTAs
 * store x[send_idx[j]] (x = that call's x) into the peers' slots and release `epoch` to their flags, the
 * tiles follow in the same grid: ONE launch exchanges and multiplies.  Same argument meaning as rsg_halo_push;
 * TILE kernel only. */
int rsg_spmv_halo_fuse(rsg_handle_t handle, int nsend, const int *send_idx, int npeers, void *const *peer_slots,
                       int *const *peer_flags, int epoch);

/* Same as rsg_spmv_halo_fuse; peer q is sent only the entries j in [peer_lo[q], peer_hi[q]) of the send list (NULL: all):
 * each boundary plane goes to the neighbour that reads it, not to every reader. */
int rsg_spmv_halo_fuse_ranges(rsg_handle_t handle, int nsend, const int *send_idx, int npeers, void *const *peer_slots,
                              int *const *peer_flags, const int *peer_lo, const int *peer_hi, int epoch);

/* ---- multi-GPU row-block SpMV (SURVEY.md 8b "multi-GPU entry point", 8e) -------------------------------------------- */
/* One rank's side, everything a step needs bound ONCE (the product of /root/reference/GPU/spmv.cu:184-186 for this
 * rank's row block, plus the exchange step north_star names).  The block's CSR has its columns relabelled into
 * [local entries | pad | npeers halo slots] (halo_start = first halo column); x lives in TWO images of that layout
 * (x_images[0..1], peer-mapped), used in turn so that a rank one step ahead never overwrites a halo a neighbour still reads.
 * peer_slots / peer_flags: HOST arrays of 2*npeers DEVICE pointers, image-major ([p*npeers + q]): where peer q keeps THIS
 * rank's slot of its image-p halo (NULL: q reads nothing from this rank) and the arrival flag it awaits from this rank
 * (NULL: not a neighbour).  wait_flags[p]: this rank's own nwait flags of image p.  precision 'd' | 's' | 'm'. */
int rsg_dist_create(rsg_handle_t handle, char precision, int m, int n_ext, int nnz, int base, const int *rowptr,
                    const int *colidx, const void *values, void *const *x_images, int halo_start, int nsend,
                    const int *send_idx, int npeers, void *const *peer_slots, int *const *peer_flags, const int *peer_lo,
                    const int *peer_hi, const int *const *wait_flags, int nwait, rsg_dist_t *dist);
/* y = A_block * x.  With the arrival flags rsg_dist_create sets up: ONE kernel launch on the handle's stream -- its first CTAs
 * push this rank's boundary entries of x (image of this step) into the peers' slots and release the step number to their
 * flags, its halo tiles (and always its last tile) acquire the neighbours' flags before they read x.  After
 * rsg_dist_enable_packets / rsg_dist_enable_pull: see there.  Asynchronous. */
int rsg_dist_step(rsg_dist_t dist, void *y);
/* Packet exchange for a context made by rsg_dist_create (window-kernel blocks; RSG_UNSUPPORTED otherwise): every boundary
 * entry travels as {value, step number} in ONE 16-byte store (fp32 x: 8 bytes) and the halo tiles unpack their x window out of
 * the packets, polling the step number inside each entry -- no fence + counter + flag hop between the data and its consumer,
 * which is what bounds a product of a few microseconds (8 GPUs on config 5).  local_packets: this rank's 2 packet buffers
 * (image 0, 1), (n_ext - halo_start) packets each, ZERO-filled by the caller (the fill COMPLETE, not merely queued on another stream: the
 * call writes marks into them on the handle's stream), peer-mapped; peer_packets: 2*npeers DEVICE
 * pointers, image-major like peer_slots: where entry 0 of this rank's send list lands in peer q's buffer.  ONLY for
 * exchange patterns in which every rank also reads from every rank it sends to (stencils, structurally symmetric matrices):
 * that mutual dependency keeps a rank from running two steps ahead of a reader -- the job the arrival flags did, which are
 * no longer touched.  Every rank must have returned from this call before any rank steps; local_packets == NULL switches
 * back to the flags (a caller whose ranks did not all get RSG_OK).  side_stream != 0: the packets are pushed by a small kernel
 * of their own on a high-priority stream of the context, ordered after everything queued on the handle's stream when
 * rsg_dist_step is called (an event), instead of by the product's first CTAs: the product then issues no remote store, so its
 * completion -- the next launch's dependency -- does not wait for a flush through NVLink (measured: 2-4.5 us).  rsg_spmv_multi_create does all
 * of this by itself when the pattern allows (RSG_DIST_EXCHANGE=flags | packets | packets-fused | pull in the environment). */
int rsg_dist_enable_packets(rsg_dist_t dist, void *const *local_packets, void *const *peer_packets, int side_stream);
/* Pull exchange for a context made by rsg_dist_create (any exchange pattern; window and tile kernels): the product's first CTAs
 * FETCH this rank's halo instead of pushing its boundary -- they poll the neighbours' "x of step e is final" counters over
 * NVLink, load the entries from the owners' x images, store them into this rank's own image and raise a LOCAL flag the halo
 * tiles wait for.  The product then contains no remote store: nothing to flush through NVLink when its grid completes (measured:
 * that flush delays the next launch by 2-4 us) and no fence behind remote stores (6-10 us under load).  pull_col / pull_idx:
 * DEVICE arrays of npull entries grouped by owner (halo column the entry goes to; its index in the owner's x); owner q's
 * entries are [owner_lo[q], owner_hi[q]); peer_x: 2*npeers DEVICE pointers, image-major: peer q's x image p (NULL: nothing
 * fetched from q); peer_ready: npeers DEVICE pointers to the neighbours' counters (NULL: not a neighbour); my_ready: this
 * rank's counter; local_flags: 2 ints (image 0, 1).  Counters and flags zeroed by the caller; every rank enables before any
 * rank steps.  Measured slower than the packets here (two NVLink round trips under load before the halo tiles may start); kept for patterns that
 * are not mutual.  rsg_spmv_multi_create: RSG_DIST_EXCHANGE=pull in the environment. */
int rsg_dist_enable_pull(rsg_dist_t dist, int npull, const int *pull_col, const int *pull_idx, int npeers,
                         const void *const *peer_x, const int *const *peer_ready, const int *owner_lo, const int *owner_hi,
                         int *my_ready, int *local_flags);
void *rsg_dist_x(rsg_dist_t dist); /* DEVICE pointer of the x image the NEXT step reads (write the local entries there) */
int rsg_dist_destroy(rsg_dist_t dist);

/* Chained products x_{k+1} = alpha * A * x_k on one rank's block ("repeated products, x_{k+1} <- y_k ... the SpMV
 * epilogue stores its y rows straight into the peers' x buffers").  THREE x images rotate: step e (1, 2, ...) reads image e % 3
 * and writes its rows into the local part of image (e + 1) % 3; the tiles that own rows a neighbour reads also store them into
 * that neighbour's halo slots of image (e + 1) % 3 and the last of them releases e to the neighbour's flag of that image; step
 * e + 1 acquires flags >= e before its halo tiles copy x.  The exchange of a step has the whole product to complete in instead
 * of sitting between two products.  Arguments as rsg_dist_create with 3 images (peer_slots / peer_flags: 3 * npeers pointers,
 * image-major; wait_flags: 3).  Before step 1 the caller fills image 1 completely, local entries AND halo (rsg_dist_x), and
 * zeroes the flags.  Needs the window kernel and a rank somebody reads from: RSG_UNSUPPORTED otherwise. */
int rsg_dist_create_chain(rsg_handle_t handle, char precision, int m, int n_ext, int nnz, int base, const int *rowptr,
                          const int *colidx, const void *values, void *const *x_images, int halo_start, int nsend,
                          const int *send_idx, int npeers, void *const *peer_slots, int *const *peer_flags, const int *peer_lo,
                          const int *peer_hi, const int *const *wait_flags, int nwait, rsg_dist_t *dist);
int rsg_dist_step_chain(rsg_dist_t dist, double alpha); /* one launch: x <- alpha * A * x, boundary rows pushed by their tiles */
int rsg_dist_chain_reset(rsg_dist_t dist);              /* the next step is step 1 again (image 1 refilled, flags zeroed) */

/* A whole square CSR matrix (HOST arrays, values fp64 like the reference's struct CSR) over ngpu devices of this process:
 * row blocks (row_starts: ngpu+1 cuts, or NULL for the nnz-balanced cuts of rsg_get_row_partition), halo relabelling,
 * peer access, one handle and one rsg_dist per device, one host thread per device issuing its steps.  devices: ngpu
 * ordinals (NULL: 0..ngpu-1).  RSG_UNSUPPORTED without peer access between the devices. */
int rsg_spmv_multi_create(int ngpu, const int *devices, char precision, int m, int n, int nnz, int base,
                          const int *rowptr_host, const int *colidx_host, const double *values_host,
                          const int *row_starts, rsg_multi_t *multi);
int rsg_spmv_multi_set_x(rsg_multi_t multi, const void *x_host);            /* n entries: fp64 ('d','m') or fp32 ('s') */
int rsg_spmv_multi_step(rsg_multi_t multi, int nsteps);                     /* nsteps products, asynchronous */
int rsg_spmv_multi_time(rsg_multi_t multi, int nsteps, double *us_per_step); /* timed on the devices, slowest device */
int rsg_spmv_multi_get_y(rsg_multi_t multi, void *y_host);                  /* waits, gathers the m entries of y */
/* chained products x <- alpha * A * x, nsteps times; x_host != NULL (re)starts the chain from that vector, NULL continues it;
 * us_per_step (may be NULL): device time of the slowest rank / nsteps */
int rsg_spmv_multi_chain(rsg_multi_t multi, const void *x_host, int nsteps, double alpha, double *us_per_step);
int rsg_spmv_multi_get_x(rsg_multi_t multi, void *x_host);                  /* waits, gathers the chain's current x */
/* what: 0 ranks, 1 halo bytes all ranks receive per step, 2 rsg_spmv_algo of rank 0's last product, 3 halo slot width;
 * row_starts (may be NULL): ngpu+1 ints */
int rsg_spmv_multi_info(rsg_multi_t multi, int what, long long *value, int *row_starts);
int rsg_spmv_multi_destroy(rsg_multi_t multi);

/* ---- preconditioned conjugate gradients on the three calls above -------------------------- */
/* Solves A x = b for a symmetric positive definite CSR matrix (device arrays; x holds the initial guess on entry).
 * The reference contains no Krylov loop -- it benchmarks the kernels such a loop is made of (mixed-precision sparse
 * solvers, README.md:5-7): every iteration is one spmv_csr_d, one sptrsv_lu_* and three fused vector kernels whose
 * scalars stay on the device.  lu_precision: 0 = no preconditioner (lu_values ignored), 'd' = fp64 ILU(0) factors
 * from ilu0_factor_d, 's' = fp32 factors from ilu0_factor_s applied inside the fp64 iteration (r rounded to fp32,
 * z widened back).  Stops when |r|_2 <= rtol |b|_2 (recurrence residual) or after max_iters; *iters and *relres
 * (= |r|_2 / |b|_2) may be NULL.  Synchronous (the host reads |r|^2 once per iteration). */
int rsg_pcg_d(rsg_handle_t handle, int n, int nnz, int base, const int *rowptr, const int *colidx, const double *values,
              const void *lu_values, int lu_precision, const double *b, double *x, int max_iters, double rtol,
              int *iters, double *relres);
/* BiCGStab (van der Vorst 1992), right-preconditioned by the same ILU(0) factors, for matrices that are not
 * symmetric positive definite (config 3's FEM-like matrix has unsymmetric values).  Same arguments and stopping
 * rule; one iteration = 2 SpMV + 2 preconditioner applications + 5 fused vector kernels. */
int rsg_bicgstab_d(rsg_handle_t handle, int n, int nnz, int base, const int *rowptr, const int *colidx, const double *values,
                   const void *lu_values, int lu_precision, const double *b, double *x, int max_iters, double rtol,
                   int *iters, double *relres);

/* Restarted flexible GMRES(30) (Saad & Schultz 1986; the preconditioned basis is kept, so fp32 factors cost no
 * accuracy), right-preconditioned by the same ILU(0) factors, modified Gram-Schmidt.
 * Same arguments and stopping rule as rsg_pcg_d; an iteration is one Arnoldi step = 1 SpMV + 1 preconditioner
 * application + (j+2) fused subtract-and-dot launches at step j of the cycle; the host reads the Hessenberg column
 * once per step; *relres is the rotated right-hand side's residual estimate (the true residual is recomputed at
 * each restart).  No counterpart in the reference (README.md:5-7 names the use only); SURVEY.md 8f-2. */
int rsg_gmres_d(rsg_handle_t handle, int n, int nnz, int base, const int *rowptr, const int *colidx, const double *values,
                const void *lu_values, int lu_precision, const double *b, double *x, int max_iters, double rtol,
                int *iters, double *relres);

/* ---- host-pointer convenience wrappers (allocate, H2D, run, D2H, free; synchronous) ------- */
int spmv_csr_d_host(rsg_handle_t handle, int m, int n, int nnz, int base,
                    const int *rowptr, const int *colidx, const double *values,
                    double alpha, const double *x, double beta, double *y);
int spmv_csr_s_host(rsg_handle_t handle, int m, int n, int nnz, int base,
                    const int *rowptr, const int *colidx, const float *values,
                    float alpha, const float *x, float beta, float *y);
int spmv_csr_mixed_host(rsg_handle_t handle, int m, int n, int nnz, int base,
                        const int *rowptr, const int *colidx, const float *values,
                        double alpha, const double *x, double beta, double *y);
int ilu0_factor_d_host(rsg_handle_t handle, int n, int nnz, int base,
                       const int *rowptr, const int *colidx, double *values_inout, int *zero_pivot_row);
int ilu0_factor_s_host(rsg_handle_t handle, int n, int nnz, int base,
                       const int *rowptr, const int *colidx, float *values_inout, int *zero_pivot_row);
int sptrsv_lu_d_host(rsg_handle_t handle, int n, int nnz, int base,
                     const int *rowptr, const int *colidx, const double *lu_values,
                     const double *b, double *y);
int sptrsv_lu_s_host(rsg_handle_t handle, int n, int nnz, int base,
                     const int *rowptr, const int *colidx, const float *lu_values,
                     const float *b, float *y);

/* ---- device-resident matrix for repeated host-vector products (x in, y out per call) ------ */
/* precision: 'd' (fp64), 's' (fp32: values cast on upload, like spmv.cu:63-67), 'm' (mixed).
 * values_host is always fp64, as in the reference CSR struct. */
int rsg_matrix_upload(rsg_handle_t handle, char precision, int m, int n, int nnz, int base,
                      const int *rowptr_host, const int *colidx_host, const double *values_host,
                      rsg_matrix_t *matrix);
/* x_host / y_host are fp64 for 'd' and 'm', fp32 for 's' (passed as void*).  Synchronous.  With beta == 0 the
 * call is pipelined over row chunks fixed at upload time: chunk c's product starts as soon as the prefix of x
 * its rows reference has arrived, and its slice of y goes back to the host while chunk c+1 is multiplied, so on a
 * banded matrix the two PCIe directions and the kernel overlap (a matrix whose first rows reference all of x
 * degrades to "x in, then product and y out overlapped").  Page-locked caller buffers are used in place. */
int rsg_matrix_spmv_host(rsg_handle_t handle, rsg_matrix_t matrix, double alpha,
                         const void *x_host, double beta, void *y_host);
int rsg_matrix_free(rsg_handle_t handle, rsg_matrix_t matrix);

#ifdef __cplusplus
}
#endif
#endif /* REASPASOL_GPU_H */
