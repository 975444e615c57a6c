// EXPERIMENT (tools/build_variants.sh persist; needs -DMARCH_NB_TABLE -DMARCH_PERSIST): the marching
// sweep as a persistent kernel.  One warp per CTA as in march_kernel, but a CTA keeps taking work
// items from a device counter and rolls its two-stage TMA pipeline straight from the post-roll
// stage of one item into the pre-roll stage of the next: no drain, no CTA start-up chain (items ->
// neighbour row -> first copy -> first data, two to four DRAM latencies of ~1-3 us against an item
// lifetime of 40-80 us), no CTA launch.  The per-item context is one row of the plan's neighbour
// table (plan.cu build_nbtab): entry 15 of row 0 carries nk | valid1 << 8, so items[] is not read.
// Arithmetic, stage layout and copies are march_kernel's (march_steps.cuh, march_dev.cuh): outputs
// are bitwise the same.
#pragma once
#include "march_steps.cuh"

#if !defined(MARCH_NB_TABLE)
#error "MARCH_PERSIST needs MARCH_NB_TABLE"
#endif

namespace {

template <int R, bool CUBE, int VAR, int P>
__global__ void __launch_bounds__(32)
    __maxnreg__(CUBE ? MARCH_CUBE_REGS : (R == 1 ? 128 : (R == 2 ? 168 : MARCH_STAR4_REGS)))
        march_persist_kernel(const __grid_constant__ MarchParams prm) {
  using L = dev::StageLayout<R, CUBE, P>;
  constexpr int H = 8 / P;
  extern __shared__ __align__(1024) char smem[];
  const int lane = threadIdx.x;

  const uint32_t smem0 = dev::smem_u32(smem);
  if ((smem0 & 511u) != 0) __trap();
  const uint32_t bar0 = smem0 + 2 * L::BYTES;
  if (lane == 0) {
    dev::mbar_init(bar0, 1);
    dev::mbar_init(bar0 + 8, 1);
    dev::fence_barrier_init();
    dev::fence_proxy_async();
  }
  __syncwarp();

  dev::LaneK<R, CUBE, P> lk;
  lk.init(lane);
  const dev::PieceK pk = dev::piece_of<R, CUBE, P>(lane < 12 ? lane : 0);
  constexpr int MAIN_PLANE_BYTES = 2 * L::C_PL + 2 * L::I_PL + L::JSUM_PL + (CUBE ? 4 * L::X_PL : 0);
  constexpr int ROLL_PLANE_BYTES = CUBE ? MAIN_PLANE_BYTES : 2 * L::C_PL;
  const int lane_off = lk.j * 8 + lk.ih * 4;

  /* blockIdx.y = field, as in march_kernel; every field has its own counter of items */
  const FieldDev &fd = prm.fields[blockIdx.y];
  double *const out_base = fd.out;
  const unsigned long long out_stride = fd.out_step;
  unsigned *const counter = prm.counter + blockIdx.y;
  auto take = [&]() -> unsigned {
    unsigned w = 0;
    if (lane == 0) w = atomicAdd(counter, 1u);
    return __shfl_sync(0xffffffffu, w, 0);
  };
  /* row `row` of work w's neighbour table: lanes 0..11 piece neighbours, 12/13 own bricks,
   * 14 pairing flags, 15 (row 0 only) nk | valid1 << 8 */
  auto row_of = [&](unsigned w, int row) -> unsigned {
    const unsigned *nbt = prm.nbtab + (size_t)w * prm.nbtab_stride;
    return lane < 16 ? __ldg(nbt + row * 16 + lane) : 0u;
  };
  /* copies of one stage into buffer g & 1 */
  auto issue = [&](uint32_t g, bool roll, int ck, unsigned nb) {
    const uint32_t bar = bar0 + 8 * (g & 1);
    const uint32_t dst = smem0 + (g & 1) * L::BYTES + pk.off;
    if (lane == 0) dev::mbar_expect_tx(bar, roll ? R * ROLL_PLANE_BYTES : P * MAIN_PLANE_BYTES);
    const int np = roll ? L::NPIECE_ROLL : L::NPIECE_MAIN;
    const unsigned flags = __shfl_sync(0xffffffffu, nb, 14); /* 0 in the roll rows */
    const bool lead = lane < 12 && ((flags >> (16 + (lane & 15))) & 1u);
    const bool follow = lane < 12 && ((flags >> (lane & 15)) & 1u);
    if (lane < np && !follow) {
      const CUtensorMap *map =
          roll ? &fd.maps.roll[pk.map] : (lead ? &fd.maps.pair[pk.pairmap] : &fd.maps.main[pk.map]);
      dev::tma_load_4d(dst, map, pk.ci, pk.cj, ck, (int)nb, bar);
    }
  };

  unsigned w = take();
  if (w >= prm.nitems) return;
  uint32_t g = 0; /* stages processed by this warp so far: buffer g & 1, barrier phase (g >> 1) & 1 */
  unsigned nbn = row_of(w, 0);
  issue(g, true, 8 - R, nbn);
  unsigned hdr = __shfl_sync(0xffffffffu, nbn, 15);
  nbn = row_of(w, 1);

  for (;;) {
    const int nk = (int)(hdr & 0xffu), nsteps = nk * H + 2;
    const bool valid1 = ((hdr >> 8) & 1u) != 0;
    const unsigned wn = take(); /* fetched a whole item ahead */
    const bool more = wn < prm.nitems;
    unsigned hdr_next = 0;

    double carry[2 * R][4];
#pragma unroll
    for (int a = 0; a < 2 * R; ++a)
#pragma unroll
      for (int m = 0; m < 4; ++m) carry[a][m] = 0.0;
    dev::OutPtrs o;
    o.cur = o.prev = nullptr;
    o.curv = o.prevv = false;
    unsigned bcur = 0, bnext = 0; /* the lane's own brick in the layer of stage s / s+1 */

    for (int s = 0; s < nsteps; ++s, ++g) {
      bcur = bnext;
      if (s + 1 < nsteps) {
        dev::fence_proxy_async();
        const bool roll = s + 1 == nsteps - 1;
        issue(g + 1, roll, roll ? 0 : (s % H) * P, nbn);
        bnext = __shfl_sync(0xffffffffu, nbn, 12 + lk.bb);
        if (s + 2 < nsteps)
          nbn = row_of(w, s + 2 == nsteps - 1 ? nk + 1 : 1 + (s + 1) / H);
        else if (more)
          nbn = row_of(wn, 0);
      } else if (more) {
        /* last stage of this item: the next item's pre-roll goes into the other buffer now */
        dev::fence_proxy_async();
        issue(g + 1, true, 8 - R, nbn);
        hdr_next = __shfl_sync(0xffffffffu, nbn, 15);
        nbn = row_of(wn, 1);
      }
      dev::mbar_wait(bar0 + 8 * (g & 1), (g >> 1) & 1);
      const char *st = smem + (g & 1) * L::BYTES;
      int mode = 1, zbase = 0;
      if (s == 0) {
        mode = 0;
      } else if (s == nsteps - 1) {
        mode = 2;
        zbase = 8;
      } else {
        const int layer = (s - 1) / H, half = (s - 1) % H;
        zbase = half * P;
        if (half == 0) {
          o.prev = o.cur;
          o.prevv = o.curv && layer > 0;
          o.cur = out_base + (size_t)bcur * out_stride + lane_off;
          o.curv = (lk.bb == 0) || valid1;
        }
      }
      if constexpr (CUBE && R == 1) {
        if (mode == 1)
          dev::cube27_step<P, P, 1>(st, lk, prm.c, carry, o, zbase);
        else if (mode == 0)
          dev::cube27_step<P, R, 0>(st, lk, prm.c, carry, o, zbase);
        else
          dev::cube27_step<P, R, 2>(st, lk, prm.c, carry, o, zbase);
      } else if constexpr (CUBE) {
        if (mode == 1)
          dev::cube_step<P, P, 1>(st, lk, prm.c, carry, o, zbase);
        else if (mode == 0)
          dev::cube_step<P, R, 0>(st, lk, prm.c, carry, o, zbase);
        else
          dev::cube_step<P, R, 2>(st, lk, prm.c, carry, o, zbase);
      } else {
        if (mode == 1)
          dev::star_step<R, VAR, P, P, 1>(st, lk, prm.c, carry, o, zbase);
        else if (mode == 0)
          dev::star_step<R, VAR, P, R, 0>(st, lk, prm.c, carry, o, zbase);
        else
          dev::star_step<R, VAR, P, R, 2>(st, lk, prm.c, carry, o, zbase);
      }
      __syncwarp();
    }
    if (!more) break;
    w = wn;
    hdr = hdr_next;
  }
}

/* resident CTAs of the kernel on this device (cached), grid = min(work, that) */
template <int R, bool CUBE, int VAR, int P>
int launch_persist(bricks_plan *plan, int stencil_id, const FieldDev *fields, int nfields, const double *c,
                   cudaStream_t s) {
  using L = dev::StageLayout<R, CUBE, P>;
  const int smem = 2 * L::BYTES + 64;
  auto kern = march_persist_kernel<R, CUBE, VAR, P>;
  static int resident[64] = {0};
  int &cap = resident[plan->ctx->device & 63];
  if (!cap) {
    BRK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    BRK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
    int per_sm = 0;
    BRK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, 32, smem));
    cap = per_sm * plan->ctx->sm_count;
    if (cap <= 0) {
      bricks::set_error("march_persist_kernel does not fit on an SM");
      return -6;
    }
  }
  MarchParams prm;
  int nitems = 0;
  const unsigned maxcol = bricks::column_length(stencil_id);
  if (int rc = bricks::plan_items(plan, maxcol, &prm.items, &nitems)) return rc;
  if (int rc = bricks::plan_nbtab(plan, maxcol, &prm.nbtab, &prm.nbtab_stride)) return rc;
  if (plan->ncounters < nfields) {
    BRK_CUDA(cudaStreamSynchronize(s)); /* an earlier launch may still be reading the old array */
    BRK_CUDA(cudaFree(plan->counter_dev));
    plan->counter_dev = nullptr;
    plan->ncounters = 0;
    BRK_CUDA(cudaMalloc(&plan->counter_dev, sizeof(unsigned) * nfields));
    plan->ncounters = nfields;
  }
  BRK_CUDA(cudaMemsetAsync(plan->counter_dev, 0, sizeof(unsigned) * nfields, s));
  prm.grid = plan->grid_dev;
  prm.adj = plan->adj_dev;
  prm.fields = fields;
  prm.gd0 = plan->gdims[0];
  prm.gd1 = plan->gdims[1];
  for (int i = 0; i < 27; ++i) prm.c[i] = c[i];
  prm.counter = plan->counter_dev;
  prm.nitems = (unsigned)nitems;
  /* the resident CTAs are split evenly over the fields */
  unsigned gx = (unsigned)cap / (unsigned)nfields;
  if (gx < 1) gx = 1;
  if (gx > (unsigned)nitems) gx = (unsigned)nitems;
  kern<<<dim3(gx, nfields), 32, smem, s>>>(prm);
  BRK_CUDA(cudaGetLastError());
  ++bricks::g_launches;
  return 0;
}

}  // namespace
