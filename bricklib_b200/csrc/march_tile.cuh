// The "tile" form of the 125-point cube sweep (sm_100a): one warp marches a column of brick
// layers along k for a footprint of 2 x 2 bricks (16 x 16 points per plane), and every lane owns a
// 2 (j) x 4 (i) register tile.  Why: the cube is bound by shared-memory traffic, not by HBM or the
// fp64 pipe -- the 1 x 4 tile of the window form (march_window.cuh) reads 5 rows x 8 columns = 10
// doubles per point and plane from shared memory; the 2 x 4 tile reads 6 rows x 8 columns for two
// output rows = 6 doubles per point, and the 16 x 16 footprint cuts the halo a plane drags in from
// 1.875x to 1.56x of its own points.
//
// Stage layout: per k-plane one padded image, rows j = -2..17, columns i = -2..17 (ten 16-byte
// chunks per row), stored as TWO sub-images -- even image rows, then odd image rows -- each with
// a row pitch of 176 B.  A lane reads image rows 2*rp .. 2*rp+5, i.e. rows rp, rp+1, rp+2 of each
// sub-image: the eight lanes of an LDS.128 phase (four quads x two adjacent row pairs) are then one
// sub-image row apart (176 = 48 mod 128) and hit eight distinct 16-byte bank groups, and every load
// address is ONE lane base plus an immediate.  The image is filled by per-lane 16-byte cp.async
// (LDGSTS) from the sixteen bricks around the footprint that BrickInfo::adj names
// (include/brick.h:96-127); arithmetic is cube_scatter (march_steps.cuh), shared with the other
// forms, so results are bitwise the same.
#pragma once
#include "march_steps.cuh"

namespace dev {

template <int P>
struct TileLayout {
  static constexpr int SUBROWS = 10, NCHUNK = 10;
  static constexpr int PITCH = 176;
  /* one parity sub-image, padded to 64 mod 128 so the two 64-byte halves of a source line (an even
   * and an odd image row) land on disjoint banks */
#ifdef MARCH_TILE_NOPAD /* 2 x 4 planes = 28 160 B per warp: eight warps fill the SM's 228 KB exactly */
  static constexpr int SUB = SUBROWS * PITCH;
#else
  static constexpr int SUB = SUBROWS * PITCH + 96;
#endif
  static constexpr int PLANE = 2 * SUB;
  static constexpr int BYTES = P * PLANE;
};

/* Producer side.  A stage is filled by 16-byte cp.async whose lanes walk memory in its own order, so
 * that one instruction touches as few 128-byte lines as possible (the SM's L1TEX queue serves one
 * line of a scattered LDGSTS every ~2 cycles and is shared with every LDS): per plane
 *   4 x centre brick   lane = 16-byte chunk of the brick's plane (512 contiguous bytes, 4 lines)
 *   1 x j-faces        8 lanes each: rows 6,7 of the two bricks above, rows 0,1 of the two below
 *   1 x i-faces        8 lanes each: last chunk column of the two bricks left, first of the two right
 * and per stage one instruction for the 2x2-element corners of every plane.  The sixteen source
 * bricks of a stage are fetched from BrickInfo::adj by lanes 0..15 (one entry each) a stage ahead. */
struct TileProducer {
  int nb_base, nb_slot; /* lane t < 16: which column brick (0..3 = b00,b10,b01,b11) and which adj slot */
  uint32_t c_dst;       /* centre: image offset of (row 2 + lane/4, chunk 1 + lane%4) */
  int j_src, i_src, x_src;          /* byte offsets inside the source brick (plane 0) */
  uint32_t j_dst, i_dst, x_dst;     /* image offsets (plane 0) */
  int j_from, i_from, x_from;       /* lane that holds the source brick id */
  bool x_valid;

  template <int P>
  static __device__ __forceinline__ uint32_t img(int rr, int cc) {
    return (uint32_t)((rr & 1) * TileLayout<P>::SUB + (rr >> 1) * TileLayout<P>::PITCH + cc * 16);
  }
  template <int P>
  __device__ __forceinline__ void init(int lane) {
    /* adjacency entry of lane t: 0-3 centres, 4-5 above, 6-7 below, 8-9 left, 10-11 right,
     * 12-15 corners TL TR BL BR; slot = (dj+1)*3 + (di+1) within the dk block */
    const int t = lane & 15;
    /* bases {0,1,2,3, 0,1,2,3, 0,2,1,3, 0,1,2,3}, slots {4,4,4,4, 1,1,7,7, 3,3,5,5, 0,2,6,8}, packed */
    nb_base = (int)((0xe4d8e4e4u >> (2 * t)) & 3u);
    nb_slot = (int)((0x8620553377114444ull >> (4 * t)) & 15ull);
    c_dst = img<P>(2 + (lane >> 2), 1 + (lane & 3));
    const int grp = lane >> 3, l8 = lane & 7;
    /* j-faces: grp 0,1 = above brick column 0,1 (rows 6,7); grp 2,3 = below (rows 0,1) */
    j_from = 4 + grp;
    j_src = (grp < 2 ? 6 * 64 : 0) + l8 * 16;
    j_dst = img<P>((grp < 2 ? 0 : 18) + (l8 >> 2), 1 + 4 * (grp & 1) + (l8 & 3));
    /* i-faces: grp 0,1 = left of brick row 0,1 (columns 6,7); grp 2,3 = right (columns 0,1) */
    i_from = 8 + grp;
    i_src = l8 * 64 + (grp < 2 ? 48 : 0);
    i_dst = img<P>(2 + 8 * (grp & 1) + l8, grp < 2 ? 0 : 9);
    /* corners: lane = 8*plane + 2*corner + row */
    const int kk = lane >> 3, corner = (lane >> 1) & 3, row = lane & 1;
    x_valid = kk < P;
    x_from = 12 + corner;
    x_src = kk * 512 + (corner < 2 ? 6 + row : row) * 64 + ((corner & 1) ? 0 : 48);
    x_dst = (uint32_t)(kk * TileLayout<P>::PLANE) + img<P>((corner < 2 ? 0 : 18) + row, (corner & 1) ? 9 : 0);
  }
};

template <int P, int NZ, int MODE>
__device__ __forceinline__ void cube_tile_step(const char *st, const double *__restrict__ c,
                                               double (&carryA)[4][4], double (&carryB)[4][4],
                                               const OutPtrs &oa, int zbase) {
  using T = TileLayout<P>;
  constexpr int R = 2;
  OutPtrs ob = oa; /* second row of the tile: eight doubles further */
  ob.cur = oa.cur + 8;
  ob.prev = oa.prev + 8;
  double accA[NZ + 4][4], accB[NZ + 4][4];
#pragma unroll
  for (int a = 0; a < NZ + 4; ++a)
#pragma unroll
    for (int m = 0; m < 4; ++m) {
      accA[a][m] = a < 4 ? carryA[a][m] : 0.0;
      accB[a][m] = a < 4 ? carryB[a][m] : 0.0;
    }

#pragma unroll
  for (int z = 0; z < NZ; ++z) {
    double eA[3][8], eB[3][8];
    {
      double w[6][8];
#pragma unroll
      for (int a = 0; a < 6; ++a)
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const double2 v = *reinterpret_cast<const double2 *>(st + z * T::PLANE + (a & 1) * T::SUB +
                                                               (a >> 1) * T::PITCH + q * 16);
          w[a][2 * q] = v.x;
          w[a][2 * q + 1] = v.y;
        }
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        eA[0][q] = w[2][q];
        eA[1][q] = w[1][q] + w[3][q];
        eA[2][q] = w[0][q] + w[4][q];
        eB[0][q] = w[3][q];
        eB[1][q] = w[2][q] + w[4][q];
        eB[2][q] = w[1][q] + w[5][q];
      }
    }
    cube_scatter<NZ, MODE>(eA, c, accA, z);
    cube_scatter<NZ, MODE>(eB, c, accB, z);
    store_done_nb<R, MODE>(oa, zbase, z, accA[z]);
    store_done_nb<R, MODE>(ob, zbase, z, accB[z]);
  }
#pragma unroll
  for (int a = 0; a < 4; ++a)
#pragma unroll
    for (int m = 0; m < 4; ++m) {
      carryA[a][m] = accA[NZ + a][m];
      carryB[a][m] = accB[NZ + a][m];
    }
}

}  // namespace dev

#ifndef MARCH_TILE_REGS
#define MARCH_TILE_REGS 255
#endif
#ifndef MARCH_TILE_STAGES
#define MARCH_TILE_STAGES 3
#endif

namespace {
/* 125-point cube, tile form (march_tile.cuh): 2x2-brick footprint, 2x4 points per lane */
template <int P, int NS>
__global__ void __maxnreg__(MARCH_TILE_REGS) cube_tile_kernel(const __grid_constant__ MarchParams prm) {
  using T = dev::TileLayout<P>;
  constexpr int R = 2, H = 8 / P;
  extern __shared__ __align__(1024) char smem[];
  const int lane = threadIdx.x;
  const MarchItem it = prm.items[blockIdx.x];
  const FieldDev &fd = prm.fields[blockIdx.y];
  const int nk = it.nk, nsteps = nk * H + 2;
  const uint32_t smem0 = dev::smem_u32(smem);

  /* brick ids of the four columns of the footprint: lane t holds layer t */
  unsigned c00 = 0, c10 = 0, c01 = 0, c11 = 0;
  if (lane < nk) {
    const size_t cell = ((size_t)(it.gk0 + lane) * prm.gd1 + it.gj) * prm.gd0 + it.gi;
    c00 = prm.grid[cell];
    c10 = it.valid1 ? prm.grid[cell + 1] : prm.adj[(size_t)c00 * 27 + 14];
    c01 = it.validj ? prm.grid[cell + prm.gd0] : prm.adj[(size_t)c00 * 27 + 16];
    c11 = (it.valid1 && it.validj) ? prm.grid[cell + prm.gd0 + 1] : prm.adj[(size_t)c01 * 27 + 14];
  }

  dev::TileProducer tp = {};
  tp.template init<P>(lane);
  const char *inb = reinterpret_cast<const char *>(fd.in);
  const size_t brick_bytes = (size_t)fd.in_step * 8;

  /* lane t < 16 fetches source brick t of stage s (march_tile.cuh: TileProducer) */
  auto load_nb = [&](int s) -> unsigned {
    const int dk = (s == 0) ? -1 : ((s == nsteps - 1) ? +1 : 0);
    const int layer = (s == 0) ? 0 : ((s == nsteps - 1) ? nk - 1 : (s - 1) / H);
    const unsigned b00 = __shfl_sync(0xffffffffu, c00, layer), b10 = __shfl_sync(0xffffffffu, c10, layer);
    const unsigned b01 = __shfl_sync(0xffffffffu, c01, layer), b11 = __shfl_sync(0xffffffffu, c11, layer);
    const unsigned base = tp.nb_base < 2 ? (tp.nb_base == 0 ? b00 : b10) : (tp.nb_base == 2 ? b01 : b11);
    return dev::ld_nc_pinned(prm.adj + (size_t)base * 27 + (dk + 1) * 9 + tp.nb_slot);
  };
  auto issue = [&](int s, unsigned nb) {
    const bool roll = (s == 0) || (s == nsteps - 1);
    const int ck = (s == 0) ? 8 - R : (roll ? 0 : ((s - 1) % H) * P);
    const int npl = roll ? R : P;
    const uint32_t stage = smem0 + (s % NS) * T::BYTES;
    const char *src0 = inb + (size_t)(ck * 512);
    const char *pc[4];
#pragma unroll
    for (int b = 0; b < 4; ++b)
      pc[b] = src0 + ((size_t)__shfl_sync(0xffffffffu, nb, b) * brick_bytes + (size_t)(lane * 16));
    const char *pj = src0 + ((size_t)__shfl_sync(0xffffffffu, nb, tp.j_from) * brick_bytes + (size_t)tp.j_src);
    const char *pi = src0 + ((size_t)__shfl_sync(0xffffffffu, nb, tp.i_from) * brick_bytes + (size_t)tp.i_src);
    const char *px = src0 + ((size_t)__shfl_sync(0xffffffffu, nb, tp.x_from) * brick_bytes + (size_t)tp.x_src);
#pragma unroll
    for (int k = 0; k < P; ++k) {
      if (k < npl) {
#pragma unroll
        for (int b = 0; b < 4; ++b) /* brick b = (row b>>1, column b&1): 4 sub-image rows, 4 chunks on */
          dev::cp_async16_nb(stage + tp.c_dst + k * T::PLANE + (b >> 1) * 4 * T::PITCH + (b & 1) * 64, pc[b] + k * 512);
        dev::cp_async16_nb(stage + tp.j_dst + k * T::PLANE, pj + k * 512);
        dev::cp_async16_nb(stage + tp.i_dst + k * T::PLANE, pi + k * 512);
      }
    }
    if (tp.x_valid && (lane >> 3) < npl) dev::cp_async16_nb(stage + tp.x_dst, px);
    dev::cp_async_commit();
  };

  double carryA[2 * R][4], carryB[2 * R][4];
#pragma unroll
  for (int a = 0; a < 2 * R; ++a)
#pragma unroll
    for (int m = 0; m < 4; ++m) carryA[a][m] = carryB[a][m] = 0.0;

  /* consumer lane = 4*rp + quad: rows 2rp, 2rp+1, columns 4*quad .. 4*quad+3 of the footprint */
  const int rp = lane >> 2, quad = lane & 3;
  const int cbase = rp * T::PITCH + quad * 32;
  const int lane_off = ((2 * rp) & 7) * 8 + (quad & 1) * 4;
  const bool lane_valid = ((quad < 2) || it.valid1) && ((rp < 4) || it.validj);
  dev::OutPtrs o;
  o.cur = o.prev = nullptr;
  o.curv = o.prevv = false;

  /* NS-deep pipeline, one cp.async group per loop trip (empty near the end of the column), so the
   * group of stage s is always the NS-th newest when stage s is consumed */
  unsigned nbn;
#pragma unroll
  for (int t = 0; t < NS - 1; ++t) {
    nbn = load_nb(t);
    issue(t, nbn);
  }
  nbn = load_nb(NS - 1 < nsteps ? NS - 1 : 0);

  for (int s = 0; s < nsteps; ++s) {
    dev::cp_async_wait<NS - 2>(); /* this lane's chunks of stage s have landed ... */
    __syncwarp();                 /* ... and so have every other lane's */
    /* refill the buffer the previous trip finished with; the copies carry no compiler barrier, so
     * their address arithmetic interleaves with this stage's loads and arithmetic */
    if (s + NS - 1 < nsteps) {
      issue(s + NS - 1, nbn);
      if (s + NS < nsteps) nbn = load_nb(s + NS);
    } else {
      dev::cp_async_commit();
    }
    const char *st = smem + (s % NS) * T::BYTES + cbase;
    int mode = 1, zbase = 0;
    if (s == 0) {
      mode = 0;
    } else if (s == nsteps - 1) {
      mode = 2;
      zbase = 8;
    } else {
      const int layer = (s - 1) / H, part = (s - 1) % H;
      zbase = part * P;
      if (part == 0) {
        const unsigned b00 = __shfl_sync(0xffffffffu, c00, layer), b10 = __shfl_sync(0xffffffffu, c10, layer);
        const unsigned b01 = __shfl_sync(0xffffffffu, c01, layer), b11 = __shfl_sync(0xffffffffu, c11, layer);
        const unsigned mine = rp < 4 ? (quad < 2 ? b00 : b10) : (quad < 2 ? b01 : b11);
        o.prev = o.cur;
        o.prevv = o.curv && layer > 0;
        o.cur = fd.out + (size_t)mine * fd.out_step + lane_off;
        o.curv = lane_valid;
      }
    }
    if (mode == 1)
      dev::cube_tile_step<P, P, 1>(st, prm.c, carryA, carryB, o, zbase);
    else if (mode == 0)
      dev::cube_tile_step<P, R, 0>(st, prm.c, carryA, carryB, o, zbase);
    else
      dev::cube_tile_step<P, R, 2>(st, prm.c, carryA, carryB, o, zbase);
    __syncwarp(); /* every lane is done reading this stage before it is refilled */
  }
}

template <int P, int NS>
int launch_tile(bricks_plan *plan, int stencil_id, const FieldDev *fields, int nfields, const double *c, cudaStream_t s) {
  using T = dev::TileLayout<P>;
  const int smem = NS * T::BYTES;
  auto kern = cube_tile_kernel<P, NS>;
  static bool configured[64] = {false};
  bool &done = configured[plan->ctx->device & 63];
  if (!done) {
    BRK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    BRK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
    done = true;
  }
  MarchParams prm;
  int nitems = 0;
  if (int rc = bricks::plan_items(plan, bricks::column_length(stencil_id), &prm.items, &nitems, 2)) return rc;
  prm.grid = plan->grid_dev;
  prm.adj = plan->adj_dev;
  prm.fields = fields;
  prm.gd0 = plan->gdims[0];
  prm.gd1 = plan->gdims[1];
  for (int i = 0; i < 10; ++i) prm.c[i] = c[i];
  kern<<<dim3(nitems, nfields), 32, smem, s>>>(prm);
  BRK_CUDA(cudaGetLastError());
  ++bricks::g_launches;
  return 0;
}

}  // namespace
