// The "window" form of the 125-point cube sweep (sm_100a).  Same work decomposition as
// march_kernel (one warp marches a column of brick layers for two bricks adjacent in i, lane =
// four consecutive points of one row, k-neighbours in register accumulators), but the stage in
// shared memory is ONE padded plane image per k -- rows j = -2..9, columns i = -2..17 of the
// two-brick footprint, cut from the nine bricks BrickInfo::adj names (include/brick.h:96-127) --
// filled by per-lane 16-byte cp.async (SASS LDGSTS) instead of one TMA box per piece:
//   * the SM's TMA engine accepts a copy only every ~46 cycles (profiles/r01_tma_issue_probe.txt),
//     which made a 9-12 box stage the floor of the TMA version; here a stage is 4*P LDGSTS per lane
//     whose addresses are one of three per-lane bases plus an immediate;
//   * every lane reads its 8-wide window (quad + 2 columns either side) of rows j-2..j+2 straight
//     from the image: no shuffles, no edge-lane selects, no per-row address tables -- one base
//     register and immediates -- which is where the TMA version spent its registers.
// Row pitch 176 B (= 48 mod 128) makes the eight lanes of an LDS.128 phase (two rows x four quads)
// hit eight distinct 16-byte bank groups.  Arithmetic is cube_scatter (march_steps.cuh), shared
// with the TMA version, so both give bitwise the same result.
#pragma once
#include "march_steps.cuh"

namespace dev {

template <int P>
struct WinLayout {
  static constexpr int NROW = 12, NCHUNK = 10; /* rows -2..9; 16-byte chunks = columns -2..17 */
  static constexpr int PITCH = 176;
  static constexpr int PLANE = NROW * PITCH;
  static constexpr int BYTES = P * PLANE;
};

/* What the lane copies, constant over the march.  Lane = 10*r3 + cc (lanes 30, 31 idle): chunk
 * column cc of image rows 3g + r3, g = 0..3, for every plane of the stage. */
struct WinProducer {
  bool valid, useb1;
  int di;         /* brick column of the chunk relative to its pair brick */
  int offA, offB, offC; /* byte offset inside the source brick for g = 0, 1 (g = 2: +192), 3 */
  bool a_low, c_high;   /* g = 0 reads the -j neighbour, g = 3 the +j neighbour */
  uint32_t dst0;

  __device__ __forceinline__ void init(int lane) {
    const int r3 = lane / 10, cc = lane - 10 * r3;
    valid = lane < 30;
    useb1 = cc >= 5;
    di = cc == 0 ? -1 : (cc == 9 ? +1 : 0);
    const int i_in = cc == 0 ? 6 : (cc == 9 ? 0 : 2 * ((cc - 1) & 3));
    a_low = r3 < 2;
    c_high = r3 > 0;
    offA = (a_low ? r3 + 6 : 0) * 64 + i_in * 8;
    offB = (1 + r3) * 64 + i_in * 8;
    offC = (c_high ? r3 - 1 : 7) * 64 + i_in * 8;
    dst0 = (uint32_t)(r3 * WinLayout<4>::PITCH + cc * 16);
  }
};

struct WinNb {
  unsigned m, z, p; /* bricks at dj = -1, 0, +1 in the lane's brick column */
};

template <int P, int NZ, int MODE>
__device__ __forceinline__ void cube_window_step(const char *st, const double *__restrict__ c,
                                                 double (&carry)[4][4], const OutPtrs &o, int zbase) {
  using W = WinLayout<P>;
  constexpr int R = 2;
  double acc[NZ + 4][4];
#pragma unroll
  for (int a = 0; a < NZ + 4; ++a)
#pragma unroll
    for (int m = 0; m < 4; ++m) acc[a][m] = a < 4 ? carry[a][m] : 0.0;

#pragma unroll
  for (int z = 0; z < NZ; ++z) {
    double e[3][8];
    {
      double w[5][8];
#pragma unroll
      for (int a = 0; a < 5; ++a)
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const double2 v = *reinterpret_cast<const double2 *>(st + (z * W::NROW + a) * W::PITCH + q * 16);
          w[a][2 * q] = v.x;
          w[a][2 * q + 1] = v.y;
        }
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        e[0][q] = w[2][q];
        e[1][q] = w[1][q] + w[3][q];
        e[2][q] = w[0][q] + w[4][q];
      }
    }
    cube_scatter<NZ, MODE>(e, c, acc, z);
    store_done<R, MODE>(o, zbase, z, acc[z]);
  }
#pragma unroll
  for (int a = 0; a < 4; ++a)
#pragma unroll
    for (int m = 0; m < 4; ++m) carry[a][m] = acc[NZ + a][m];
}

}  // namespace dev

#ifndef MARCH_WINDOW_REGS
#define MARCH_WINDOW_REGS 168
#endif
#ifndef MARCH_WINDOW_STAGES
#define MARCH_WINDOW_STAGES 2
#endif

namespace {
/* 125-point cube, window form (march_window.cuh): padded plane images filled by per-lane cp.async */
template <int P, int NS>
__global__ void __maxnreg__(MARCH_WINDOW_REGS) cube_window_kernel(const __grid_constant__ MarchParams prm) {
  using W = dev::WinLayout<P>;
  constexpr int R = 2, H = 8 / P;
  extern __shared__ __align__(1024) char smem[];
  const int lane = threadIdx.x;
  const MarchItem it = prm.items[blockIdx.x];
  const FieldDev &fd = prm.fields[blockIdx.y];
  const int nk = it.nk, nsteps = nk * H + 2;
  const uint32_t smem0 = dev::smem_u32(smem);

  unsigned col0 = 0, col1 = 0; /* brick ids of the column: lane t holds layer t */
  if (lane < nk) {
    const size_t cell = ((size_t)(it.gk0 + lane) * prm.gd1 + it.gj) * prm.gd0 + it.gi;
    col0 = prm.grid[cell];
    col1 = it.valid1 ? prm.grid[cell + 1] : prm.adj[(size_t)col0 * 27 + 14];
  }

  dev::WinProducer wp;
  wp.init(lane);
  const char *inb = reinterpret_cast<const char *>(fd.in);
  const size_t brick_bytes = (size_t)fd.in_step * 8;

  /* the three bricks (dj = -1, 0, +1) of the lane's brick column for stage s */
  auto load_nb = [&](int s) -> dev::WinNb {
    const int dk = (s == 0) ? -1 : ((s == nsteps - 1) ? +1 : 0);
    const int layer = (s == 0) ? 0 : ((s == nsteps - 1) ? nk - 1 : (s - 1) / H);
    const unsigned b0 = __shfl_sync(0xffffffffu, col0, layer);
    const unsigned b1 = __shfl_sync(0xffffffffu, col1, layer);
    const unsigned *row = prm.adj + (size_t)(wp.useb1 ? b1 : b0) * 27 + (dk + 1) * 9 + (wp.di + 1);
    dev::WinNb nb;
    nb.m = __ldg(row);
    nb.z = __ldg(row + 3);
    nb.p = __ldg(row + 6);
    return nb;
  };
  auto issue = [&](int s, const dev::WinNb &nb) {
    const bool roll = (s == 0) || (s == nsteps - 1);
    const int ck = (s == 0) ? 8 - R : (roll ? 0 : ((s - 1) % H) * P);
    const int npl = roll ? R : P;
    const char *pa = inb + ((size_t)(wp.a_low ? nb.m : nb.z) * brick_bytes + (size_t)(wp.offA + ck * 512));
    const char *pb = inb + ((size_t)nb.z * brick_bytes + (size_t)(wp.offB + ck * 512));
    const char *pc = inb + ((size_t)(wp.c_high ? nb.p : nb.z) * brick_bytes + (size_t)(wp.offC + ck * 512));
    const uint32_t dst = smem0 + (s % NS) * W::BYTES + wp.dst0;
    if (wp.valid) {
#pragma unroll
      for (int k = 0; k < P; ++k) {
        if (k < npl) {
          dev::cp_async16(dst + (k * W::NROW + 0) * W::PITCH, pa + k * 512);
          dev::cp_async16(dst + (k * W::NROW + 3) * W::PITCH, pb + k * 512);
          dev::cp_async16(dst + (k * W::NROW + 6) * W::PITCH, pb + k * 512 + 192);
          dev::cp_async16(dst + (k * W::NROW + 9) * W::PITCH, pc + k * 512);
        }
      }
    }
    dev::cp_async_commit();
  };

  double carry[2 * R][4];
#pragma unroll
  for (int a = 0; a < 2 * R; ++a)
#pragma unroll
    for (int m = 0; m < 4; ++m) carry[a][m] = 0.0;

  /* consumer lane = 4*j + 2*bb + ih: the eight lanes of an LDS.128 phase cover two rows x four quads */
  const int cj = lane >> 2, cbb = (lane >> 1) & 1, cih = lane & 1;
  const int cbase = cj * W::PITCH + cbb * 64 + cih * 32;
  const int lane_off = cj * 8 + cih * 4;
  dev::OutPtrs o;
  o.cur = o.prev = nullptr;
  o.curv = o.prevv = false;

  /* NS-deep pipeline, one cp.async group per loop trip (empty near the end of the column), so the
   * group of stage s is always the NS-th newest when stage s is consumed */
  dev::WinNb nbn;
#pragma unroll
  for (int t = 0; t < NS - 1; ++t) {
    nbn = load_nb(t < nsteps ? t : 0);
    if (t < nsteps) issue(t, nbn);
    else dev::cp_async_commit();
  }
  nbn = load_nb(NS - 1 < nsteps ? NS - 1 : 0);

  for (int s = 0; s < nsteps; ++s) {
    if (s + NS - 1 < nsteps) {
      issue(s + NS - 1, nbn);
      if (s + NS < nsteps) nbn = load_nb(s + NS);
    } else {
      dev::cp_async_commit();
    }
    dev::cp_async_wait<NS - 1>(); /* this lane's chunks of stage s have landed ... */
    __syncwarp();                 /* ... and so have every other lane's */
    const char *st = smem + (s % NS) * W::BYTES + cbase;
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
        const unsigned b0 = __shfl_sync(0xffffffffu, col0, layer);
        const unsigned b1 = __shfl_sync(0xffffffffu, col1, layer);
        o.prev = o.cur;
        o.prevv = o.curv && layer > 0;
        o.cur = fd.out + (size_t)(cbb ? b1 : b0) * fd.out_step + lane_off;
        o.curv = (cbb == 0) || (it.valid1 != 0);
      }
    }
    if (mode == 1)
      dev::cube_window_step<P, P, 1>(st, prm.c, carry, o, zbase);
    else if (mode == 0)
      dev::cube_window_step<P, R, 0>(st, prm.c, carry, o, zbase);
    else
      dev::cube_window_step<P, R, 2>(st, prm.c, carry, o, zbase);
    __syncwarp(); /* every lane is done reading this stage before it is refilled */
  }
}

template <int P>
int launch_window(bricks_plan *plan, int stencil_id, const FieldDev *fields, int nfields, const double *c, cudaStream_t s) {
  using W = dev::WinLayout<P>;
  const int smem = MARCH_WINDOW_STAGES * W::BYTES;
  auto kern = cube_window_kernel<P, MARCH_WINDOW_STAGES>;
  static bool configured[64] = {false};
  bool &done = configured[plan->ctx->device & 63];
  if (!done) {
    BRK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    BRK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
    done = true;
  }
  MarchParams prm;
  int nitems = 0;
  if (int rc = bricks::plan_items(plan, bricks::column_length(stencil_id), &prm.items, &nitems)) return rc;
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
