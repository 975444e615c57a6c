// The 4-D path: the 9-point star of stencils/mpi9pt.py on 4x4x4x4 fp64 bricks (weak/main-4d.cpp, a
// CPU-only driver in the reference: brick_stencil :62-75 runs the generated kernel over every brick of
// the ghost-including grid, `skip / TILE` bricks inside on each side).
//
// A brick is 256 doubles = 2 KiB, row-major [l][k][j][i]; a row of four doubles is one 32-byte sector.
// Every brick needs its own 64 rows plus one 4x4x4 face of each of its eight face neighbours, which come
// through the 81-slot adjacency list exactly as in the reference (include/brick.h:229-246: slot
// ((dl*3 + dk)*3 + dj)*3 + di, digit 1 = same brick), so out-of-grid neighbours resolve to brick 0.
//
// line_kernel (the product path): one thread per row position (l,k,j); the 64 threads of a brick march
// along the contiguous grid axis.  The i-neighbours of a row are then the row this thread held one step
// ago and the row it loads for the next step: the i-faces -- the only faces that would use a quarter of
// every 32-byte sector -- cost no memory traffic at all.  The other six neighbour rows are whole sectors
// read through L1; a CTA covers 2 x 2 lines adjacent in j and k so that half of those rows are L1 hits.
// HBM-bound: 16 algorithmic bytes per point (8 read + 8 written), no shared memory, no barrier.
// gather_kernel (verification twin): one thread per point straight through the accessor arithmetic.
// The bodies live in stencil4d_body.cuh, which also compiles as host code for a thread-by-thread emulation.
#include <cstdlib>
#include <cstring>

#include "common.cuh"
#include "stencil4d_body.cuh"

struct bricks_plan4d {
  bricks_ctx *ctx = nullptr;
  unsigned gdims[4] = {0, 0, 0, 0};
  unsigned nbricks = 0;
  unsigned *grid_dev = nullptr;
  unsigned *adj_dev = nullptr;
};

namespace {

using s4::kNAdj;
using s4::Sweep4;

/* LJ x LK lines per CTA; at 80 registers 768 threads fit an SM whatever the CTA size */
template <int LJ, int LK, bool WIDE>
__global__ void __launch_bounds__(64 * LJ * LK, 768 / (64 * LJ * LK)) line_kernel(const __grid_constant__ Sweep4 p) {
  s4::line_body<LJ, LK, WIDE>(p, blockIdx.x, threadIdx.x);
}

struct LineVariant {
  const char *name; /* "<LJ><LK><wide>" */
  int lj, lk;
  void (*fn)(const Sweep4);
};
const LineVariant kLineVariants[] = {
    {"220", 2, 2, line_kernel<2, 2, false>}, {"221", 2, 2, line_kernel<2, 2, true>},
    {"410", 4, 1, line_kernel<4, 1, false>}, {"411", 4, 1, line_kernel<4, 1, true>},
    {"210", 2, 1, line_kernel<2, 1, false>}, {"211", 2, 1, line_kernel<2, 1, true>},
    {"110", 1, 1, line_kernel<1, 1, false>}, {"111", 1, 1, line_kernel<1, 1, true>},
};
constexpr const char *kDefaultLineVariant = "221"; /* profiles/r03_4d/variants.log */

/* BRICKS_4D_VARIANT=<LJ><LK><wide> picks another CTA shape / load width (experiments and tests) */
const LineVariant *line_variant() {
  static const LineVariant *chosen = [] {
    const char *env = getenv("BRICKS_4D_VARIANT");
    const char *want = env && *env ? env : kDefaultLineVariant;
    for (const LineVariant &v : kLineVariants)
      if (!strcmp(v.name, want)) return &v;
    return (const LineVariant *)nullptr;
  }();
  return chosen;
}

__global__ void __launch_bounds__(256) gather_kernel(const __grid_constant__ Sweep4 p) {
  s4::gather_body(p, blockIdx.x, threadIdx.x);
}

int fill_params(bricks_plan4d *plan, const double *in, double *out, size_t in_step, size_t out_step,
                unsigned skip_bricks, int stream, Sweep4 *p, bool *empty) {
  BRK_ARG(plan && in && out && in != out && stream >= 0 && stream < bricks::kNumStreams);
  /* rows are read as two 16-byte halves and written as one 32-byte store */
  BRK_ARG(((uintptr_t)in % 32) == 0 && ((uintptr_t)out % 32) == 0 && in_step % 4 == 0 && out_step % 4 == 0);
  BRK_ARG(in_step >= 256 && out_step >= 256);
  *empty = false;
  for (int a = 0; a < 4; ++a)
    if (plan->gdims[a] <= 2 * skip_bricks) *empty = true;
  p->grid = plan->grid_dev;
  p->adj = plan->adj_dev;
  p->in = in;
  p->out = out;
  p->in_step = in_step;
  p->out_step = out_step;
  p->n0 = plan->gdims[0], p->n1 = plan->gdims[1], p->n2 = plan->gdims[2], p->n3 = plan->gdims[3];
  p->s = skip_bricks;
  p->njg = p->nkg = 0;
  return 0;
}

}  // namespace

extern "C" {

int bricks_plan4d_create(bricks_ctx *ctx, const unsigned *grid_host, const unsigned gdims[4],
                         const unsigned *adj_host, unsigned nbricks, bricks_plan4d **out) {
  BRK_ARG(ctx && grid_host && gdims && adj_host && out && nbricks > 0);
  size_t cells = 1;
  for (int a = 0; a < 4; ++a) {
    BRK_ARG(gdims[a] > 0 && gdims[a] < 65536);
    cells *= gdims[a];
  }
  BRK_ARG(cells < (1ull << 31));
  for (size_t c = 0; c < cells; ++c) {
    const unsigned b = grid_host[c];
    if (b >= nbricks) {
      bricks::set_error("bricks_plan4d_create: grid cell %zu names brick %u of %u", c, b, nbricks);
      return -1;
    }
    const unsigned *row = adj_host + (size_t)b * kNAdj;
    /* the reference reaches the centre through slot 40 as well (include/brick.h:229-246) */
    if (row[40] != b) {
      bricks::set_error("bricks_plan4d_create: adj[%u][40] = %u is not the brick itself", b, row[40]);
      return -1;
    }
    for (int n = 0; n < kNAdj; ++n)
      if (row[n] >= nbricks) {
        bricks::set_error("bricks_plan4d_create: adj[%u][%d] = %u is outside the storage (%u bricks)", b, n, row[n],
                          nbricks);
        return -1;
      }
  }
  BRK_CUDA(cudaSetDevice(ctx->device));
  bricks_plan4d *pl = new bricks_plan4d;
  pl->ctx = ctx;
  pl->nbricks = nbricks;
  for (int a = 0; a < 4; ++a) pl->gdims[a] = gdims[a];
  cudaError_t e = cudaMalloc(&pl->grid_dev, cells * sizeof(unsigned));
  if (e == cudaSuccess) e = cudaMalloc(&pl->adj_dev, (size_t)nbricks * kNAdj * sizeof(unsigned));
  if (e == cudaSuccess) e = cudaMemcpy(pl->grid_dev, grid_host, cells * sizeof(unsigned), cudaMemcpyHostToDevice);
  if (e == cudaSuccess)
    e = cudaMemcpy(pl->adj_dev, adj_host, (size_t)nbricks * kNAdj * sizeof(unsigned), cudaMemcpyHostToDevice);
  if (e != cudaSuccess) {
    bricks::set_error("bricks_plan4d_create: %s", cudaGetErrorString(e));
    cudaFree(pl->grid_dev);
    cudaFree(pl->adj_dev);
    delete pl;
    return (int)e;
  }
  *out = pl;
  return 0;
}

int bricks_plan4d_destroy(bricks_plan4d *plan) {
  if (!plan) return 0;
  cudaSetDevice(plan->ctx->device);
  cudaFree(plan->grid_dev);
  cudaFree(plan->adj_dev);
  delete plan;
  return 0;
}

int bricks_stencil4d(bricks_plan4d *plan, const double *in, double *out, size_t in_step, size_t out_step,
                     unsigned skip_bricks, int stream) {
  Sweep4 p;
  bool empty;
  if (int rc = fill_params(plan, in, out, in_step, out_step, skip_bricks, stream, &p, &empty)) return rc;
  if (empty) return 0;
  BRK_CUDA(cudaSetDevice(plan->ctx->device));
  const LineVariant *v = line_variant();
  if (!v) {
    bricks::set_error("BRICKS_4D_VARIANT=%s is not a built variant", getenv("BRICKS_4D_VARIANT"));
    return -1;
  }
  const unsigned long long nblocks = s4::line_blocks(&p, v->lj, v->lk);
  BRK_ARG(nblocks < (1ull << 31));
  static std::atomic<unsigned> configured{0}; /* bit per device: all of L1 as cache (no shared memory used) */
  const unsigned bit = 1u << (plan->ctx->device & 31);
  if (!(configured.load() & bit)) {
    BRK_CUDA(cudaFuncSetAttribute(v->fn, cudaFuncAttributePreferredSharedMemoryCarveout, 0));
    configured.fetch_or(bit);
  }
  v->fn<<<(unsigned)nblocks, 64 * v->lj * v->lk, 0, plan->ctx->streams[stream]>>>(p);
  BRK_CUDA(cudaGetLastError());
  ++bricks::g_launches;
  return 0;
}

int bricks_stencil4d_naive(bricks_plan4d *plan, const double *in, double *out, size_t in_step, size_t out_step,
                           unsigned skip_bricks, int stream) {
  Sweep4 p;
  bool empty;
  if (int rc = fill_params(plan, in, out, in_step, out_step, skip_bricks, stream, &p, &empty)) return rc;
  if (empty) return 0;
  BRK_CUDA(cudaSetDevice(plan->ctx->device));
  const unsigned long long nblocks = s4::gather_blocks(p);
  BRK_ARG(nblocks < (1ull << 31));
  gather_kernel<<<(unsigned)nblocks, 256, 0, plan->ctx->streams[stream]>>>(p);
  BRK_CUDA(cudaGetLastError());
  ++bricks::g_launches;
  return 0;
}

}  // extern "C"
