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

#ifdef __cplusplus
}
#endif
#endif /* BRICKS_B200_H */
