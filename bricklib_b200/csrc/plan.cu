// Launch plans: the host side of the reference's `<<<dim3(strideb0,strideb1,strideb2),32>>>`
// launch shape (weak/main.cu:277-281, stencils/3axis.cu:163-169, strong/main.cu:464-475),
// re-expressed as marching work items, plus TMA tensor-map construction over a brick field.
#include <algorithm>
#include <cstring>

#include <cstdlib>

#include "common.cuh"

namespace bricks {

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *,
                                  const cuuint64_t *, const cuuint64_t *, const cuuint32_t *,
                                  const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn lookup_encode() {
  void *p = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
      q == cudaDriverEntryPointSuccess)
    return (EncodeTiledFn)p;
  return nullptr;
}
static EncodeTiledFn get_encode() {
  static const EncodeTiledFn fn = lookup_encode(); /* initialised once, thread-safe (C++11) */
  return fn;
}

/* A field is viewed as the 4-D tensor [brick][k][j][i] of fp64 with strides
 * (step*8, 512, 64, 8) bytes: the layout contract of include/brick.h:229-246 for fold <4,8>. */
int encode_map(CUtensorMap *m, const void *base, unsigned nbricks, size_t step_elems, unsigned bi,
               unsigned bj, unsigned bk, bool swizzle64, unsigned bb, unsigned elem_bytes) {
  EncodeTiledFn fn = get_encode();
  if (!fn) {
    set_error("cuTensorMapEncodeTiled is not available from this driver");
    return -3;
  }
  cuuint64_t dims[4] = {8, 8, 8, nbricks};
  /* fp64: rows of 64 bytes, planes of 512; fp32 bricks (bElem = float): 32 and 256 */
  cuuint64_t strides[3] = {8ull * elem_bytes, 64ull * elem_bytes, (cuuint64_t)step_elems * elem_bytes};
  cuuint32_t box[4] = {bi, bj, bk, bb};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = fn(m, elem_bytes == 8 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT64 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, const_cast<void *>(base), dims, strides, box,
                  estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  swizzle64 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_NONE,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled(box %ux%ux%u, step %zu) failed with CUresult %d", bi, bj, bk,
              step_elems, (int)r);
    return 2000 + (int)r;
  }
  return 0;
}

/* longest column a single warp marches; chosen so a 66-layer column splits evenly and the
 * pre/post-roll overhead stays below ~5 % */
unsigned column_length(int stencil_id) {
#ifdef BRICKS_EXPERIMENT /* experiment builds only (tools/build_variants.sh): BRICKS_MAXCOL overrides */
  static const int forced = [] {
    const char *e = getenv("BRICKS_MAXCOL");
    return (e && atoi(e) > 0 && atoi(e) < 32) ? atoi(e) : 0;
  }();
  if (forced) return (unsigned)forced;
#endif
  /* short columns give the scheduler more, smaller items (bandwidth-bound stars); long columns
   * amortise the pre-/post-roll planes (arithmetic-bound cube) */
  switch (stencil_id) {
    case BRICKS_ST_125PT: return 17;
    case BRICKS_ST_27PT_COEFF: return 11;
    case BRICKS_ST_25PT: return 10;
    case BRICKS_ST_13PT: return 9;
    default: return 7;
  }
}

static void build_items(const unsigned gdims[3], int nboxes, const unsigned *boxes, unsigned kMaxColumn,
                        std::vector<MarchItem> &items, std::vector<unsigned> *cells) {
  for (int b = 0; b < nboxes; ++b) {
    const unsigned *lo = boxes + 6 * b, *hi = lo + 3;
    if (hi[0] <= lo[0] || hi[1] <= lo[1] || hi[2] <= lo[2]) continue;
    unsigned nk = hi[2] - lo[2];
    unsigned nseg = (nk + kMaxColumn - 1) / kMaxColumn;
    unsigned seglen = (nk + nseg - 1) / nseg;
    for (unsigned s = 0; s < nseg; ++s) {
      unsigned k0 = lo[2] + s * seglen;
      unsigned k1 = std::min(hi[2], k0 + seglen);
      if (k0 >= k1) continue;
      for (unsigned j = lo[1]; j < hi[1]; ++j)
        for (unsigned i = lo[0]; i < hi[0]; i += 2) {
          MarchItem it;
          it.gi = (unsigned short)i;
          it.gj = (unsigned short)j;
          it.gk0 = (unsigned short)k0;
          it.nk = (unsigned short)(k1 - k0);
          it.valid1 = (i + 1 < hi[0]) ? 1u : 0u;
          it.pad_ = 0u;
          items.push_back(it);
        }
    }
    if (!cells) continue;
    for (unsigned k = lo[2]; k < hi[2]; ++k)
      for (unsigned j = lo[1]; j < hi[1]; ++j)
        for (unsigned i = lo[0]; i < hi[0]; ++i)
          cells->push_back((k * gdims[1] + j) * gdims[0] + i);
  }
}


int plan_items(bricks_plan *p, unsigned maxcol, const MarchItem **items_dev, int *nitems) {
  const unsigned key = maxcol;
  auto it = p->items.find(key);
  if (it == p->items.end()) {
    std::vector<MarchItem> items;
    build_items(p->gdims, (int)(p->boxes.size() / 6), p->boxes.data(), maxcol, items, nullptr);
    bricks_plan::ItemList l;
    l.n = (int)items.size();
    cudaStream_t s = p->ctx->streams[BRICKS_STREAM_COMPUTE];
    BRK_CUDA(cudaMalloc(&l.dev, items.size() * sizeof(MarchItem)));
    BRK_CUDA(cudaMemcpyAsync(l.dev, items.data(), items.size() * sizeof(MarchItem), cudaMemcpyHostToDevice, s));
    BRK_CUDA(cudaStreamSynchronize(s)); /* `items` dies here */
    it = p->items.emplace(key, l).first;
  }
  *items_dev = it->second.dev;
  *nitems = it->second.n;
  return 0;
}

}  // namespace bricks

using namespace bricks;

namespace {

/* boxes inside the grid, brick lists and the first item list of a plan over `grid_host` */
int plan_fill(bricks_plan *p, const unsigned *grid_host, int nboxes, const unsigned *boxes) {
  for (int b = 0; b < nboxes; ++b)
    for (int d = 0; d < 3; ++d) BRK_ARG(boxes[6 * b + 3 + d] <= p->gdims[d]);
  std::vector<MarchItem> items;
  std::vector<unsigned> cells;
  build_items(p->gdims, nboxes, boxes, 11, items, &cells);
  BRK_ARG(!items.empty());
  p->boxes.assign(boxes, boxes + 6 * (size_t)nboxes);
  std::vector<unsigned> blist(cells.size());
  for (size_t n = 0; n < cells.size(); ++n) blist[n] = grid_host[cells[n]];
  cudaStream_t s = p->ctx->streams[BRICKS_STREAM_COMPUTE];
  BRK_CUDA(cudaMalloc(&p->blist_dev, blist.size() * sizeof(unsigned)));
  BRK_CUDA(cudaMemcpyAsync(p->blist_dev, blist.data(), blist.size() * sizeof(unsigned),
                           cudaMemcpyHostToDevice, s));
  BRK_CUDA(cudaStreamSynchronize(s)); /* host vectors die here */
  p->nlist = (long)blist.size();
  return 0;
}

struct PlanGuard { /* releases the plan on every early return; handed to the caller at the end */
  bricks_plan *p;
  ~PlanGuard() { bricks_plan_destroy(p); }
};

}  // namespace

extern "C" {

int bricks_plan_create(bricks_ctx *ctx, const unsigned *grid_host, const unsigned gdims[3],
                       const unsigned *adj_host, unsigned nbricks, int nboxes, const unsigned *boxes,
                       bricks_plan **out) {
  BRK_ARG(ctx && grid_host && gdims && adj_host && boxes && out && nboxes > 0 && nbricks > 0);
  BRK_ARG(gdims[0] < 65536 && gdims[1] < 65536 && gdims[2] < 65536);
  BRK_CUDA(cudaSetDevice(ctx->device));
  PlanGuard guard{new bricks_plan};
  bricks_plan *p = guard.p;
  p->ctx = ctx;
  p->nbricks = nbricks;
  size_t ncell = 1;
  for (int d = 0; d < 3; ++d) {
    p->gdims[d] = gdims[d];
    ncell *= gdims[d];
  }
  for (size_t c = 0; c < ncell; ++c) BRK_ARG(grid_host[c] < nbricks);
  /* every neighbour a kernel may follow names a brick of the storage (the BrickDecomp sentinel is
   * brick 0, include/brick-mpi.h:353) */
  for (size_t n = 0; n < (size_t)nbricks * 27; ++n) BRK_ARG(adj_host[n] < nbricks);
  /* The marching kernels take the +i brick of a pair and the k-neighbours of a column from the
   * GRID, every other neighbour through adj; the reference accessor resolves all of them through
   * adj (brick.h:229-246).  The two agree when BrickInfo is consistent with the grid lattice
   * (init_grid, BrickDecomp::getBrickInfo): checked here, once, so that a caller with a custom
   * adjacency (permuted or periodic-wrapped) gets an error instead of silently different results. */
  const size_t plane = (size_t)gdims[0] * gdims[1];
  for (size_t c = 0; c < ncell; ++c) {
    const unsigned b = grid_host[c];
    if ((c % gdims[0]) + 1 < gdims[0] && adj_host[(size_t)b * 27 + 14] != grid_host[c + 1]) {
      set_error("bricks_plan_create: adj[%u][14] = %u but the grid's +i neighbour is %u: BrickInfo is not "
                "consistent with the grid lattice", b, adj_host[(size_t)b * 27 + 14], grid_host[c + 1]);
      return -1;
    }
    if (c + plane < ncell && adj_host[(size_t)b * 27 + 22] != grid_host[c + plane]) {
      set_error("bricks_plan_create: adj[%u][22] = %u but the grid's +k neighbour is %u: BrickInfo is not "
                "consistent with the grid lattice", b, adj_host[(size_t)b * 27 + 22], grid_host[c + plane]);
      return -1;
    }
  }
  p->grid_host.assign(grid_host, grid_host + ncell);
  if (int rc = plan_fill(p, grid_host, nboxes, boxes)) return rc;

  cudaStream_t s = ctx->streams[BRICKS_STREAM_COMPUTE];
  BRK_CUDA(cudaMalloc(&p->grid_dev, ncell * sizeof(unsigned)));
  BRK_CUDA(cudaMalloc(&p->adj_dev, (size_t)nbricks * 27 * sizeof(unsigned)));
  BRK_CUDA(cudaMemcpyAsync(p->grid_dev, grid_host, ncell * sizeof(unsigned), cudaMemcpyHostToDevice, s));
  BRK_CUDA(cudaMemcpyAsync(p->adj_dev, adj_host, (size_t)nbricks * 27 * sizeof(unsigned),
                           cudaMemcpyHostToDevice, s));
  BRK_CUDA(cudaStreamSynchronize(s));
  guard.p = nullptr;
  *out = p;
  return 0;
}

int bricks_plan_derive(bricks_plan *parent, int nboxes, const unsigned *boxes, bricks_plan **out) {
  BRK_ARG(parent && parent->owns_tables && boxes && out && nboxes > 0);
  BRK_CUDA(cudaSetDevice(parent->ctx->device));
  PlanGuard guard{new bricks_plan};
  bricks_plan *p = guard.p;
  p->ctx = parent->ctx;
  p->nbricks = parent->nbricks;
  for (int d = 0; d < 3; ++d) p->gdims[d] = parent->gdims[d];
  p->owns_tables = false;
  p->grid_dev = parent->grid_dev;
  p->adj_dev = parent->adj_dev;
  if (int rc = plan_fill(p, parent->grid_host.data(), nboxes, boxes)) return rc;
  guard.p = nullptr;
  *out = p;
  return 0;
}

int bricks_plan_destroy(bricks_plan *p) {
  if (!p) return 0;
  cudaSetDevice(p->ctx->device);
  if (p->owns_tables) {
    cudaFree(p->grid_dev);
    cudaFree(p->adj_dev);
  }
  for (auto &kv : p->items) cudaFree(kv.second.dev);
  cudaFree(p->blist_dev);
  for (auto &kv : p->fields) cudaFree(kv.second);
  delete p;
  return 0;
}

int bricks_plan_device_tables(const bricks_plan *p, const unsigned **adj_dev, const unsigned **grid_dev) {
  BRK_ARG(p);
  if (adj_dev) *adj_dev = p->adj_dev;
  if (grid_dev) *grid_dev = p->grid_dev;
  return 0;
}

}  // extern "C"
