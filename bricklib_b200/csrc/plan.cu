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

static EncodeTiledFn get_encode() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void *p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = (EncodeTiledFn)p;
  }
  return fn;
}

/* A field is viewed as the 4-D tensor [brick][k][j][i] of fp64 with strides
 * (step*8, 512, 64, 8) bytes: the layout contract of include/brick.h:229-246 for fold <4,8>. */
int encode_map(CUtensorMap *m, const void *base, unsigned nbricks, size_t step_elems, unsigned bi,
               unsigned bj, unsigned bk, bool swizzle64, unsigned bb) {
  EncodeTiledFn fn = get_encode();
  if (!fn) {
    set_error("cuTensorMapEncodeTiled is not available from this driver");
    return -3;
  }
  cuuint64_t dims[4] = {8, 8, 8, nbricks};
  cuuint64_t strides[3] = {64, 512, (cuuint64_t)step_elems * 8};
  cuuint32_t box[4] = {bi, bj, bk, bb};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 4, const_cast<void *>(base), dims, strides, box,
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
  static int forced = -1;
  if (forced < 0) {
    const char *e = getenv("BRICKS_MAXCOL"); /* tuning knob for experiments */
    forced = (e && atoi(e) > 0 && atoi(e) < 32) ? atoi(e) : 0;
  }
  if (forced) return (unsigned)forced;
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
                        std::vector<MarchItem> &items, std::vector<unsigned> *cells, unsigned fj = 1) {
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
      for (unsigned j = lo[1]; j < hi[1]; j += fj)
        for (unsigned i = lo[0]; i < hi[0]; i += 2) {
          MarchItem it;
          it.gi = (unsigned short)i;
          it.gj = (unsigned short)j;
          it.gk0 = (unsigned short)k0;
          it.nk = (unsigned short)(k1 - k0);
          it.valid1 = (i + 1 < hi[0]) ? 1u : 0u;
          it.validj = (fj > 1 && j + 1 < hi[1]) ? 1u : 0u;
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

#ifdef MARCH_NB_TABLE
/* Experiment: everything a warp looks up through grid[] and adj[] while it marches its column,
 * resolved at plan time into one row of 16 unsigned per stage source:
 *   row 0        pre-roll  (pieces cut from the -k neighbours of layer 0)
 *   row 1 + l    layer l   (same-layer neighbours; both half-stages of the layer use it)
 *   row nk + 1   post-roll (the +k neighbours of layer nk-1)
 * entries 0..11 = neighbour brick of piece p (march_dev.cuh piece_of: p&1 picks the brick of the
 * pair, p>>1 = centre, i-face, -j face, +j face, -j corners, +j corners), 12/13 = the pair's own
 * bricks, 14 = pairing flags of the main stages (bit p: piece p arrives with piece p-1's two-brick
 * copy; bit 16+p: piece p is that two-brick copy), 15 of row 0 = nk | valid1 << 8.  The kernel reads a row as ONE coalesced 64-byte
 * load, any number of stages ahead, instead of chasing items -> grid -> adj -> adj. */
static void build_nbtab(const bricks_plan *p, const std::vector<MarchItem> &items, unsigned maxcol,
                        std::vector<unsigned> &tab, unsigned &stride) {
  stride = (maxcol + 2) * 16;
  tab.assign((size_t)items.size() * stride, 0u);
  const unsigned *grid = p->grid_host.data(), *adj = p->adj_host.data();
  for (size_t n = 0; n < items.size(); ++n) {
    const MarchItem &it = items[n];
    unsigned *base = tab.data() + n * stride;
    for (unsigned row = 0; row < (unsigned)it.nk + 2; ++row) {
      const int dk = row == 0 ? -1 : (row == (unsigned)it.nk + 1 ? +1 : 0);
      const unsigned layer = row == 0 ? 0 : (row == (unsigned)it.nk + 1 ? it.nk - 1 : row - 1);
      const size_t cell = ((size_t)(it.gk0 + layer) * p->gdims[1] + it.gj) * p->gdims[0] + it.gi;
      const unsigned b0 = grid[cell];
      const unsigned b1 = it.valid1 ? grid[cell + 1] : adj[(size_t)b0 * 27 + 14];
      unsigned *r = base + row * 16;
      for (int piece = 0; piece < 12; ++piece) {
        const int useb1 = piece & 1, grp = piece >> 1, side = useb1 ? +1 : -1;
        const int di = (grp == 1 || grp >= 4) ? side : 0;
        const int dj = (grp == 2 || grp == 4) ? -1 : ((grp == 3 || grp == 5) ? +1 : 0);
        r[piece] = adj[(size_t)(useb1 ? b1 : b0) * 27 + (dk + 1) * 9 + (dj + 1) * 3 + (di + 1)];
      }
      r[12] = b0;
      r[13] = b1;
      unsigned flags = 0;
      if (dk == 0)
        for (int piece : {0, 4, 6})
          if (r[piece + 1] == r[piece] + 1) flags |= (1u << (piece + 1)) | (1u << (16 + piece));
      r[14] = flags;
      if (row == 0) r[15] = (unsigned)it.nk | (it.valid1 ? 0x100u : 0u); /* item header (march_persist.cuh) */
    }
  }
}

int plan_nbtab(bricks_plan *p, unsigned maxcol, const unsigned **tab_dev, unsigned *stride) {
  auto it = p->items.find(maxcol);
  if (it == p->items.end()) {
    set_error("plan_nbtab: no item list for column length %u", maxcol);
    return -1;
  }
  *tab_dev = it->second.nbtab_dev;
  *stride = it->second.nbtab_stride;
  return 0;
}
#endif

int plan_items(bricks_plan *p, unsigned maxcol, const MarchItem **items_dev, int *nitems, int fj) {
  const unsigned key = maxcol | (fj > 1 ? 0x100u : 0u);
  auto it = p->items.find(key);
  if (it == p->items.end()) {
    std::vector<MarchItem> items;
    build_items(p->gdims, (int)(p->boxes.size() / 6), p->boxes.data(), maxcol, items, nullptr, fj > 1 ? 2 : 1);
    bricks_plan::ItemList l;
    l.n = (int)items.size();
    cudaStream_t s = p->ctx->streams[BRICKS_STREAM_COMPUTE];
    BRK_CUDA(cudaMalloc(&l.dev, items.size() * sizeof(MarchItem)));
    BRK_CUDA(cudaMemcpyAsync(l.dev, items.data(), items.size() * sizeof(MarchItem), cudaMemcpyHostToDevice, s));
    BRK_CUDA(cudaStreamSynchronize(s)); /* `items` dies here */
#ifdef MARCH_NB_TABLE
    if (fj <= 1) {
      std::vector<unsigned> tab;
      build_nbtab(p, items, maxcol, tab, l.nbtab_stride);
      BRK_CUDA(cudaMalloc(&l.nbtab_dev, tab.size() * sizeof(unsigned)));
      BRK_CUDA(cudaMemcpy(l.nbtab_dev, tab.data(), tab.size() * sizeof(unsigned), cudaMemcpyHostToDevice));
    }
#endif
    it = p->items.emplace(key, l).first;
  }
  *items_dev = it->second.dev;
  *nitems = it->second.n;
  return 0;
}

}  // namespace bricks

using namespace bricks;

extern "C" {

int bricks_plan_create(bricks_ctx *ctx, const unsigned *grid_host, const unsigned gdims[3],
                       const unsigned *adj_host, unsigned nbricks, int nboxes, const unsigned *boxes,
                       bricks_plan **out) {
  BRK_ARG(ctx && grid_host && gdims && adj_host && boxes && out && nboxes > 0 && nbricks > 0);
  for (int b = 0; b < nboxes; ++b)
    for (int d = 0; d < 3; ++d) BRK_ARG(boxes[6 * b + 3 + d] <= gdims[d]);
  BRK_ARG(gdims[0] < 65536 && gdims[1] < 65536 && gdims[2] < 65536);
  BRK_CUDA(cudaSetDevice(ctx->device));
  /* released on every early return below; handed to the caller at the end */
  struct Guard {
    bricks_plan *p;
    ~Guard() { bricks_plan_destroy(p); }
  } guard{new bricks_plan};
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
  std::vector<MarchItem> items;
  std::vector<unsigned> cells;
  build_items(gdims, nboxes, boxes, 11, items, &cells);
  BRK_ARG(!items.empty());
  p->boxes.assign(boxes, boxes + 6 * (size_t)nboxes);
  std::vector<unsigned> blist(cells.size());
  for (size_t n = 0; n < cells.size(); ++n) blist[n] = grid_host[cells[n]];

  cudaStream_t s = ctx->streams[BRICKS_STREAM_COMPUTE];
  BRK_CUDA(cudaMalloc(&p->grid_dev, ncell * sizeof(unsigned)));
  BRK_CUDA(cudaMalloc(&p->adj_dev, (size_t)nbricks * 27 * sizeof(unsigned)));
  BRK_CUDA(cudaMalloc(&p->blist_dev, blist.size() * sizeof(unsigned)));
  BRK_CUDA(cudaMemcpyAsync(p->grid_dev, grid_host, ncell * sizeof(unsigned), cudaMemcpyHostToDevice, s));
  BRK_CUDA(cudaMemcpyAsync(p->adj_dev, adj_host, (size_t)nbricks * 27 * sizeof(unsigned),
                           cudaMemcpyHostToDevice, s));
  BRK_CUDA(cudaMemcpyAsync(p->blist_dev, blist.data(), blist.size() * sizeof(unsigned),
                           cudaMemcpyHostToDevice, s));
  BRK_CUDA(cudaStreamSynchronize(s)); /* host vectors die here */
  p->nlist = (long)blist.size();
#ifdef MARCH_NB_TABLE
  p->grid_host.assign(grid_host, grid_host + ncell);
  p->adj_host.assign(adj_host, adj_host + (size_t)nbricks * 27);
#endif
  guard.p = nullptr;
  *out = p;
  return 0;
}

int bricks_plan_destroy(bricks_plan *p) {
  if (!p) return 0;
  cudaSetDevice(p->ctx->device);
  cudaFree(p->grid_dev);
  cudaFree(p->adj_dev);
  for (auto &kv : p->items) {
    cudaFree(kv.second.dev);
#ifdef MARCH_NB_TABLE
    cudaFree(kv.second.nbtab_dev);
#endif
  }
  cudaFree(p->blist_dev);
#ifdef MARCH_NB_TABLE
  cudaFree(p->counter_dev);
#endif
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
