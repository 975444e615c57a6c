/*
 * bricks_b200.h -- the C ABI of libbricks_b200.so (B200 / sm_100a).
 *
 * This is the drop-in boundary for bricklib's GPU stencil hot path.  The reference has no
 * runtime FFI; its seam is compile time: a `brick(...)` pragma expanded by a code generator
 * inside a __global__ (weak/main.cu:35-43), the gpu* macros / mov* helpers of
 * include/brick-cuda.h + brick-gpu.h, and BrickDecomp::exchange (include/brick-mpi.h:466-495).
 * Every entry point below names the reference interface it replaces.  Plain pointers and
 * sizes only; no C++/torch types; every function returns 0 on success or a non-zero status
 * (CUDA / NCCL / argument error) and never throws or aborts.  bricks_last_error() gives the
 * text (reference: _gpuCheck prints + exit(), brick-gpu.h:23-29).
 *
 * Threading: a context is bound to one device and may be driven by one host thread at a
 * time (reference: MPI_THREAD_SERIALIZED, weak/main.cu:48).  All compute / copy / exchange
 * calls are asynchronous on the context's streams unless stated.
 *
 * Data layout contract (include/brick.h:53-127, 229-246): a brick is 8x8x8 fp64, row-major
 * [k][j][i] (fold <4,8> / <8>); element (b,k,j,i) of a field lives at
 * dat[b*step + k*64 + j*8 + i]; `step` (elements) is 512 for one field per storage or
 * n*512 for n interleaved fields; adj[b][(dk+1)*9 + (dj+1)*3 + (di+1)] is the neighbour
 * brick index, slot 13 = b itself.
 */
#ifndef BRICKS_B200_H
#define BRICKS_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct bricks_ctx bricks_ctx;         /* one per device */
typedef struct bricks_plan bricks_plan;       /* a stencil launch plan over a brick-grid box */
typedef struct bricks_exchange bricks_exchange; /* a halo-exchange plan */
typedef struct bricks_step bricks_step;       /* the sweep schedule of one rank: exchange + ST_ITER sweeps */

/* Stencil ids.  Point sets and coefficients: stencils/7pt.py:19-26 (7 runtime coefficients
 * coeff[0..6] = centre,i+1,i-1,j+1,j-1,k+1,k-1), stencils/mpi7pt.py, mpi13pt.py, mpi25pt.py,
 * mpi125pt.py with the constants of stencils/fake.h:11-33. */
enum {
  BRICKS_ST_7PT_COEFF = 0,
  BRICKS_ST_COND = 1, /* stencils/cond.py:23-30 (d3cond / d3condcu): 7pt.py's point set and coeff[0..6] on
                         max(in, 0), result |sum| */
  BRICKS_ST_27PT_COEFF = 27, /* d3pt27, stencils/3axis.cpp:165-247: 27 runtime coefficients coeff[0..26],
                                term t at (k + t/9 - 1, j + (t/3)%3 - 1, i + t%3 - 1) except term 16, which
                                the reference reads at (k+1, j+1, i); reproduced as written */
  BRICKS_ST_7PT = 7,
  BRICKS_ST_13PT = 13,
  BRICKS_ST_25PT = 25,
  BRICKS_ST_125PT = 125
};

/* streams of a context: sweeps, halo exchange (highest priority), and two copy streams so that a
 * host-fed pipeline can upload the next field and download the last result while sweeps run */
enum { BRICKS_STREAM_COMPUTE = 0, BRICKS_STREAM_COMM = 1, BRICKS_STREAM_H2D = 2, BRICKS_STREAM_D2H = 3 };

/* ---- context, memory, copies: replaces gpuMalloc/gpuMemcpy/gpuFree (brick-cuda.h:12-24),
 * movBrickInfo / movBrickStorage (brick-gpu.h:39-99, whose `unsigned` byte counts overflow
 * above 4 GiB; sizes here are size_t), copyToDevice/copyFromDevice (stencils/cudaarray.h). */
const char *bricks_last_error(void);
int bricks_device_count(int *count);
int bricks_ctx_create(int device, bricks_ctx **ctx);
int bricks_ctx_destroy(bricks_ctx *ctx);
int bricks_ctx_device(const bricks_ctx *ctx, int *device);
int bricks_malloc(bricks_ctx *ctx, void **dptr, size_t bytes);
int bricks_free(bricks_ctx *ctx, void *dptr);
int bricks_memset(bricks_ctx *ctx, void *dptr, int value, size_t bytes);
int bricks_host_alloc(void **hptr, size_t bytes); /* pinned */
/* pinned AND placed on the NUMA node the context's device hangs off (first touch from one of that
 * node's CPUs; the calling thread keeps its affinity): with one rank per GPU and a launcher that pins
 * nothing, host<->device copies of all ranks otherwise share whatever node the pages landed on.
 * *numa_node (optional) = that node, or -1 when the box does not expose one (then == bricks_host_alloc).
 * Replaces the malloc'ed staging buffers of weak/main.cu:251-275. */
int bricks_host_alloc_local(bricks_ctx *ctx, void **hptr, size_t bytes, int *numa_node);
int bricks_host_free(void *hptr);
int bricks_h2d(bricks_ctx *ctx, void *dst_dev, const void *src_host, size_t bytes, int stream);
int bricks_d2h(bricks_ctx *ctx, void *dst_host, const void *src_dev, size_t bytes, int stream);
int bricks_d2d(bricks_ctx *ctx, void *dst_dev, const void *src_dev, size_t bytes, int stream);
int bricks_sync(bricks_ctx *ctx, int stream); /* stream < 0: whole device */
/* make `stream` wait for everything enqueued so far on `other` (event record + wait) */
int bricks_stream_wait(bricks_ctx *ctx, int stream, int other);
/* the raw cudaStream_t, for callers that share streams (e.g. torch) */
int bricks_stream_handle(bricks_ctx *ctx, int stream, void **cuda_stream);

/* ---- timers: replaces the cudaEvent pairs of stencils_cu.h:13-28 and weak/main.cu:278-286 */
int bricks_timer_start(bricks_ctx *ctx, int stream, int slot);
int bricks_timer_stop(bricks_ctx *ctx, int stream, int slot);
int bricks_timer_elapsed_ms(bricks_ctx *ctx, int slot, float *ms); /* synchronises on stop */

/* ---- stencil launch plan: replaces the launch shape of
 *   brick_kernel<<<dim3(strideb0,strideb1,strideb2),32>>>(grid, in, out, stride)  (weak/main.cu:35-43,277-281)
 *   d3pt7_brick_trans<<<dim3(NB,NB,NB),32>>>(grid, bIn, bOut, coeff)              (stencils/3axis.cu:28-37,163-169)
 *   brick_kernel<<<dim3(STRIDEB*nsub,STRIDEB,STRIDEB),32>>>(grid,strideb,barr,..) (strong/main.cu:85-99,464-469)
 * `grid_host` is the host brick-index grid [gdims[2]][gdims[1]][gdims[0]] (i fastest) and
 * `adj_host` the host adjacency list (nbricks x 27); both are copied to the device.  The plan
 * covers nboxes boxes of grid cells, box b = lo[3] then hi[3] (exclusive) in boxes[6*b..],
 * i first.  Centre bricks come from the grid; every halo comes through adj (so out-of-grid
 * neighbours resolve to brick 0 exactly as in the reference). */
int bricks_plan_create(bricks_ctx *ctx, const unsigned *grid_host, const unsigned gdims[3],
                       const unsigned *adj_host, unsigned nbricks, int nboxes,
                       const unsigned *boxes, bricks_plan **plan);
/* a plan over other boxes of the SAME grid: borrows the parent's device copies of grid / adj (the
 * reference keeps one grid / BrickInfo per subdomain, weak/main.cu:90-100); destroy it before the parent */
int bricks_plan_derive(bricks_plan *parent, int nboxes, const unsigned *boxes, bricks_plan **plan);
int bricks_plan_destroy(bricks_plan *plan);
/* device copies held by the plan (for callers that want them, e.g. the exchange) */
int bricks_plan_device_tables(const bricks_plan *plan, const unsigned **adj_dev,
                              const unsigned **grid_dev);

/* One sweep out = stencil(in) over the plan's boxes, for `nfields` independent (in,out)
 * storage pairs that share the plan's grid/adjacency (strong scaling: one pair per
 * subdomain).  in/out are DEVICE field pointers (already offset to the field inside an
 * interleaved storage); steps in elements.  coeff_host: BRICKS_ST_7PT_COEFF reads 7 doubles,
 * BRICKS_ST_COND reads 7 too, BRICKS_ST_27PT_COEFF 27, others ignore it.  Asynchronous on `stream`. */
int bricks_stencil(bricks_plan *plan, int stencil_id, int nfields, const double *const *in,
                   double *const *out, size_t in_step, size_t out_step, const double *coeff_host,
                   int stream);
/* debug / verification twin: one thread per point straight through adj (the shape of
 * d3pt7_brick, stencils/3axis.cu:12-26).  Not a fallback: the product path never calls it. */
int bricks_stencil_naive(bricks_plan *plan, int stencil_id, int nfields, const double *const *in,
                         double *const *out, size_t in_step, size_t out_step,
                         const double *coeff_host, int stream);
/* fp32 bricks: the reference's element type is overridable (`#define bElem float`, include/vecscatter.h:12-14;
 * include/bricks/brick.h honours the same macro).  A brick is then 8x8x8 floats = 2 KiB, rows of 32 bytes, same
 * [k][j][i] order, `step` still in elements; everything that moves bytes (exchange, region copies, the step
 * object's communication program) is element-size agnostic.  Same marching kernels instantiated for float
 * (quads of 16 bytes, i-face pieces four elements wide, no swizzle: four rows make one 128-byte line); the
 * coefficients are rounded to float and all arithmetic is float, as in the reference's generated code. */
int bricks_stencil_f32(bricks_plan *plan, int stencil_id, int nfields, const float *const *in,
                       float *const *out, size_t in_step, size_t out_step, const double *coeff_host,
                       int stream);
int bricks_stencil_naive_f32(bricks_plan *plan, int stencil_id, int nfields, const float *const *in,
                             float *const *out, size_t in_step, size_t out_step,
                             const double *coeff_host, int stream);
/* ---- the 4-D path: the 9-point star of stencils/mpi9pt.py on 4x4x4x4 fp64 bricks, the stencil of the
 * reference's CPU-only driver weak/main-4d.cpp.  A brick is 256 doubles = 2 KiB, row-major [l][k][j][i];
 * element (b,l,k,j,i) lives at dat[b*step + ((l*4 + k)*4 + j)*4 + i]; adj has 81 slots per brick,
 * slot ((dl*3 + dk)*3 + dj)*3 + di with digit 1 = the brick itself (include/brick.h:229-246).
 * bricks_plan4d_create replaces the host grid / BrickInfo pair of weak/main-4d.cpp:112-124: `grid_host`
 * is the brick-index grid [gdims[3]][gdims[2]][gdims[1]][gdims[0]] (contiguous axis first in gdims),
 * `adj_host` the adjacency list (nbricks x 81); both are copied to the device.  Rejected: a grid cell or
 * an adjacency entry of a grid brick outside [0, nbricks), and adj[b][40] != b. */
typedef struct bricks_plan4d bricks_plan4d;
int bricks_plan4d_create(bricks_ctx *ctx, const unsigned *grid_host, const unsigned gdims[4],
                         const unsigned *adj_host, unsigned nbricks, bricks_plan4d **plan);
int bricks_plan4d_destroy(bricks_plan4d *plan);
/* One sweep of brick_stencil (weak/main-4d.cpp:62-75): out = stencil(in) for every brick of the grid that
 * is at least `skip_bricks` cells inside it on every side (the reference passes skip / TILE); bricks
 * outside that box are not written.  Centre bricks come from the grid, every neighbour through adj.  in /
 * out: DEVICE storages, 32-byte aligned, distinct; steps in elements, multiples of 4.  Asynchronous on
 * `stream`. */
int bricks_stencil4d(bricks_plan4d *plan, const double *in, double *out, size_t in_step, size_t out_step,
                     unsigned skip_bricks, int stream);
/* verification twin (one thread per point through the accessor arithmetic); never on the product path */
int bricks_stencil4d_naive(bricks_plan4d *plan, const double *in, double *out, size_t in_step, size_t out_step,
                           unsigned skip_bricks, int stream);

/* number of kernels launched by this library so far in this process */
int bricks_launch_count(unsigned long long *count);

/* ---- region copies: replaces cudaCopy<<<nlinks,64>>>(RegionLink*) (strong/main.cu:62-83) */
typedef struct {
  double *from, *to; /* device pointers, 16-byte aligned */
  size_t len;        /* bytes, multiple of 16 */
} bricks_region_link;
/* links_host is copied to the device once per call unless links_dev_cached != NULL */
int bricks_links_upload(bricks_ctx *ctx, const bricks_region_link *links_host, int n,
                        bricks_region_link **links_dev);
int bricks_copy_regions(bricks_ctx *ctx, const bricks_region_link *links_dev, int n, int stream);
/* release a table made by bricks_links_upload (after the copies that use it have completed) */
int bricks_links_free(bricks_ctx *ctx, bricks_region_link *links_dev);
/* The same region copies for callers that may not allocate (CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * links_host (HOST memory, absolute device addresses in from / to) is staged into a table
 * that lives in a file-scope `static __device__` array of the library -- 56 KiB held by every GPU that
 * loads the library until the process exits -- and the copy kernel of bricks_copy_regions runs over it:
 * no cudaMalloc / cudaFree / allocating routine on this path.  n > BRICKS_COPY_STATIC_MAX_LINKS is
 * rejected with an error (-1; use bricks_links_upload + bricks_copy_regions for longer tables).  Every
 * call on a device shares the one table: calls on the SAME stream are ordered, calls that overlap on
 * the GPU (different streams, or host threads driving the same device) race on it.  The table keeps
 * the last call's entries past n; they are never read.  Asynchronous on `stream`. */
#define BRICKS_COPY_STATIC_MAX_LINKS 2048
int bricks_copy_regions_static(bricks_ctx *ctx, const bricks_region_link *links_host, int n, int stream);

/* ---- array <-> brick copies on the device: replaces copyToBrick<3> / copyFromBrick<3>
 * (include/bricksetup.h:172-221; cell range and strides of iter_grid, :139-160) for arrays that
 * already live in device memory.  dimlist / padding / ghost are in elements, i first, with the
 * reference's meaning: the array has dimlist + 2*(padding + ghost) elements per axis, the brick grid
 * (dimlist + 2*ghost)/8 cells per axis, and only the cells of the interior `dimlist` region are
 * copied.  grid_dev is the device brick-index grid (e.g. from bricks_plan_device_tables); `step` in
 * elements.  to_brick != 0: array -> bricks, else bricks -> array.  Asynchronous on `stream`. */
int bricks_copy_array(bricks_ctx *ctx, int to_brick, const long dimlist[3], const long padding[3],
                      const long ghost[3], double *arr_dev, const unsigned *grid_dev,
                      double *bricks_dev, size_t step, int stream);

/* ---- array-layout baseline (NOT the product path): the 7-point stencil with seven coefficients on a
 * conventional padded device array, the counterpart of d3pt7_arr / d3pt7_arr_warp
 * (stencils/3axis.cu:39-60) for an on-box brick-vs-array comparison.  ext = array extents in elements
 * (i first), [lo, hi) the box of points to compute (at least one element inside the array on every
 * side); coeff_host[0..6] = centre, i+1, i-1, j+1, j-1, k+1, k-1.  Asynchronous on `stream`. */
int bricks_array_stencil7(bricks_ctx *ctx, const double *in_dev, double *out_dev, const long ext[3],
                          const long lo[3], const long hi[3], const double *coeff_host, int stream);

/* ---- halo exchange: replaces BrickDecomp::exchange(BrickStorage&) (brick-mpi.h:466-495),
 * ExchangeView::exchange (brick-mpi.h:96-123) and the D2H / MPI / H2D staging of
 * weak/main.cu:251-275.  Region i: receive nbytes[i] at recv_off[i] of the storage from
 * rank recv_peer[i]; send nbytes[i] from send_off[i] to send_peer[i]; both sides enumerate
 * regions in the same order i (the reference's MPI tag).  Peers equal to `rank` become
 * on-device copies.  Offsets/sizes in BYTES relative to the storage base passed at run time. */
int bricks_comm_unique_id(void *id128);            /* ncclGetUniqueId; 128 bytes */
int bricks_comm_init(bricks_ctx *ctx, int rank, int nranks, const void *id128);
/* max_ctas > 0: ncclCommInitRankConfig with minCTAs = 1, maxCTAs = max_ctas, so that the send/recv
 * kernel of an exchange takes few SM slots from the sweep it overlaps; 0 = NCCL's default */
int bricks_comm_init_config(bricks_ctx *ctx, int rank, int nranks, const void *id128, int max_ctas);
int bricks_comm_destroy(bricks_ctx *ctx);
int bricks_exchange_create(bricks_ctx *ctx, int rank, int nregions, const size_t *recv_off,
                           const size_t *send_off, const size_t *nbytes, const int *recv_peer,
                           const int *send_peer, bricks_exchange **ex);
/* general form: receive and send sizes of a region may differ (either may be 0).  Used for the
 * per-peer packed messages of the strong-scaling driver (strong/main.cu:416-435): region = peer,
 * offsets point into one arena that holds every packed send and receive buffer. */
int bricks_exchange_create_v(bricks_ctx *ctx, int rank, int nregions, const size_t *recv_off,
                             const size_t *recv_bytes, const int *recv_peer, const size_t *send_off,
                             const size_t *send_bytes, const int *send_peer, bricks_exchange **ex);
int bricks_exchange_destroy(bricks_exchange *ex);
/* enqueue the whole exchange of one storage on `stream` (normally BRICKS_STREAM_COMM) */
int bricks_exchange_run(bricks_exchange *ex, double *storage_dev, int stream);

/* ---- the sweep schedule: replaces the hot loop `brick_func` of the MPI drivers (weak/main.cu:246-287,
 * strong/main.cu:397-510; CPU twins weak/main.cpp:172-213, strong/main.cpp:349-400): one exchange, then
 * ST_ITER ping-pong sweeps a -> b -> a ..., with the exchange hidden behind the part of the first sweep
 * that reads no ghost brick.  A step object holds the five box plans of that schedule over ONE device
 * copy of grid / adj (ghost_bricks = ghost depth in bricks, 1 for GZ = 8) and a communication program:
 * region-copy tables and exchange plans run in the order they were added, on the communication stream. */
enum { BRICKS_PLAN_FULL = 0, BRICKS_PLAN_INNER = 1, BRICKS_PLAN_SHELL = 2, BRICKS_PLAN_OWNED = 3, BRICKS_PLAN_OSKIN = 4 };
/* what the from / to members of a copy table (and the offsets of an exchange plan) are relative to */
enum { BRICKS_BASE_ABS = 0, BRICKS_BASE_FIELD = 1 /* a[0] of the step */, BRICKS_BASE_ARENA = 2 };
enum {
  BRICKS_STEP_OVERLAP = 1,      /* first sweep split into inner box (before the exchange lands) + shell */
  BRICKS_STEP_EARLY = 2,        /* last sweep: owned skin first, then start the NEXT step's exchange */
  BRICKS_STEP_EARLY_UNSPLIT = 4 /* with EARLY: a first sweep whose exchange was started early runs unsplit */
};
/* host only: the boxes (lo[3], hi[3] per box, i first) of plan `which` for a brick grid of gdims cells;
 * *nboxes = 0 when the subdomain has no inner box; `boxes` holds up to 36 entries */
int bricks_step_boxes(const unsigned gdims[3], unsigned ghost_bricks, int which, unsigned *boxes, int *nboxes);
int bricks_step_create(bricks_ctx *ctx, const unsigned *grid_host, const unsigned gdims[3],
                       const unsigned *adj_host, unsigned nbricks, unsigned ghost_bricks, bricks_step **step);
int bricks_step_destroy(bricks_step *step);
/* borrowed: BRICKS_PLAN_INNER / SHELL / OSKIN do not exist for subdomains without an inner box */
int bricks_step_plan(bricks_step *step, int which, bricks_plan **plan);
/* device scratch for packed messages, owned by the step */
int bricks_step_arena(bricks_step *step, size_t bytes, void **arena_dev);
int bricks_step_add_copy(bricks_step *step, const bricks_region_link *links_host, int n, int src_base, int dst_base);
/* base = BRICKS_BASE_FIELD or BRICKS_BASE_ARENA; take_ownership != 0: bricks_step_destroy destroys `ex`,
 * else it is borrowed (destroy it after the step) */
int bricks_step_add_exchange(bricks_step *step, bricks_exchange *ex, int base, int take_ownership);
/* One step over nfields storage pairs (a[f], b[f]); st_iter even, the result ends in a.  Everything is
 * enqueued asynchronously: sweeps on BRICKS_STREAM_COMPUTE, the communication program on
 * BRICKS_STREAM_COMM.  With BRICKS_STEP_EARLY the exchange of the next step is already in flight at
 * return: the next call must continue from the same a, and any other access to a must be preceded by
 * bricks_sync(ctx, -1) and bricks_step_reset. */
int bricks_step_run(bricks_step *step, int stencil_id, int st_iter, int nfields, double *const *a,
                    double *const *b, const double *coeff_host, int flags);
/* the communication program alone (after everything enqueued on the compute stream so far) */
int bricks_step_exchange(bricks_step *step, double *field0);
/* one sweep over one of the step's plans */
int bricks_step_sweep(bricks_step *step, int which, int stencil_id, int nfields, const double *const *in,
                      double *const *out, const double *coeff_host, int stream);
/* Copy-engine exchange: the region messages of a step as cudaMemcpyAsync over NVLink straight into the
 * peers' ghost bricks (peer-mapped memory; CUDA IPC between processes, direct pointers between threads of
 * one process), ordered by flag words and stream memory operations -- no SM-resident send/recv kernel
 * competing with the sweep the exchange hides behind (replaces the CUDA-aware MPI path of
 * weak/main.cu:270-275 the way NVLink wants it).  Every rank exports a 256-byte descriptor of the storage it
 * exchanges (field0 = the a[0] its steps run on, the start of a bricks_malloc allocation), the host layer
 * gathers the descriptors of all ranks (MPI_Allgather's role; torch.distributed in bench.py, shared memory
 * in the drivers) and hands them to connect.  Afterwards the exchanges of that storage take this path;
 * exchanges of any other storage, and packed exchanges, stay on NCCL.  All ranks must hold the same region
 * tables (equal subdomains; checked).  Call connect on every rank before the first step; destroy the step
 * only after all ranks have synchronised. */
int bricks_step_p2p_export(bricks_step *step, double *field0, int rank, void *desc256);
int bricks_step_p2p_connect(bricks_step *step, int nranks, const void *descs /* nranks x 256 bytes, by rank */);
/* back to NCCL for every exchange of the step (unmaps the peers).  Setting the path up is all-or-nothing
 * across ranks: when export or connect fails on ANY rank (CUDA IPC disabled in a container, no peer access),
 * every rank must call this -- bench.py and bricklib_b200/weak.py agree on it with one more allgather. */
int bricks_step_p2p_disable(bricks_step *step);
/* Timeline of a step, for profiles: with tracing on, bricks_step_run puts a CUDA event pair around every
 * operation it enqueues (replaces the per-phase cudaEvent / MPI_Wtime pairs of weak/main.cu:251-286).
 * bricks_step_trace_read synchronises the device and returns, for the LAST traced step, one record per
 * operation: code = BRICKS_TRACE_SWEEP + 10 * sweep + plan, BRICKS_TRACE_COPY + op or
 * BRICKS_TRACE_EXCHANGE + op (op = index in the communication program), and its start / end in ms
 * relative to the earliest start.  An event on a stream completes when the work before it has, so
 * a record's span includes time its stream spent waiting for the other one. */
enum { BRICKS_TRACE_SWEEP = 1000, BRICKS_TRACE_COPY = 2000, BRICKS_TRACE_EXCHANGE = 3000 };
int bricks_step_trace(bricks_step *step, int enable);
int bricks_step_trace_read(bricks_step *step, int cap, int *n, int *codes, float *start_ms, float *end_ms);
/* forget an exchange started early (the caller rewrote the field) */
int bricks_step_reset(bricks_step *step);

#ifdef __cplusplus
}
#endif
#endif /* BRICKS_B200_H */
