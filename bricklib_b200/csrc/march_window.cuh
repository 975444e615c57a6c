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
