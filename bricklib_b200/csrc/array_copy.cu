// Array <-> brick copies on the device.
//   bricks_copy_array   replaces copyToBrick<3> / copyFromBrick<3> (include/bricksetup.h:172-221,
//                       iteration order of iter_grid, :139-160) for arrays that already live in
//                       device memory: an application holding a conventional padded array adopts
//                       the brick path without a host round trip (the reference copies on the host
//                       and then moves the whole storage, stencils/3axis.cu:101-121).
// HBM-bound, 16 B per point (one 8-byte read, one 8-byte write).  One CTA moves a strip of
// eight bricks adjacent in i: a (k, j) row of the strip is 512 contiguous bytes of the array (one
// warp, 16 bytes per lane = four full lines) and eight 64-byte rows on the brick side (two full
// sectors each), so both sides move whole sectors.
#include "common.cuh"

namespace {

constexpr int kStrip = 8;      /* bricks per CTA along i */
constexpr int kThreads = 256;  /* 8 warps = 8 (k, j) rows per pass */

struct ArrayCopyParams {
  double *arr;
  double *bricks;
  const unsigned *grid;
  long sa1, sa2;      /* array strides of j and k (elements); i is contiguous */
  long sb1, sb2;      /* grid strides of j and k (cells) */
  long off[3];        /* array offset of brick cell 0: padding (elements) */
  unsigned lo[3], n[3]; /* first brick cell and count per axis */
  size_t step;
};

/* VEC = 2: 16-byte accesses (array rows 16-byte aligned); VEC = 1: 8-byte accesses */
template <int VEC, bool TO_BRICK>
__global__ void __launch_bounds__(kThreads) array_copy_kernel(const ArrayCopyParams p) {
  const unsigned bj = p.lo[1] + blockIdx.y, bk = p.lo[2] + blockIdx.z;
  const unsigned strip0 = blockIdx.x * kStrip; /* relative to lo[0] */
  constexpr int PER_ROW = kStrip * 8 / VEC;    /* accesses per (k, j) row of the strip */
  constexpr int ROWS = kThreads / PER_ROW;     /* rows per pass */
  const int t = threadIdx.x;
  const int row_in_pass = t / PER_ROW, a = t % PER_ROW;
  const unsigned sb = (unsigned)(a * VEC) >> 3, i = (unsigned)(a * VEC) & 7u; /* brick in strip, element */
  if (strip0 + sb >= p.n[0]) return;
  const unsigned bi = p.lo[0] + strip0 + sb;
  const unsigned b = p.grid[(size_t)bk * p.sb2 + (size_t)bj * p.sb1 + bi];
  double *bdat = p.bricks + (size_t)b * p.step + i;
  double *adat = p.arr + (p.off[2] + (long)bk * 8) * p.sa2 + (p.off[1] + (long)bj * 8) * p.sa1 + p.off[0] + (long)bi * 8 + i;
#pragma unroll 4
  for (int r = row_in_pass; r < 64; r += ROWS) {
    const int k = r >> 3, j = r & 7;
    double *ap = adat + k * p.sa2 + j * p.sa1;
    double *bp = bdat + r * 8;
    if (VEC == 2) {
      if (TO_BRICK) __stcs(reinterpret_cast<double2 *>(bp), __ldcs(reinterpret_cast<const double2 *>(ap)));
      else __stcs(reinterpret_cast<double2 *>(ap), __ldcs(reinterpret_cast<const double2 *>(bp)));
    } else {
      if (TO_BRICK) __stcs(bp, __ldcs(ap));
      else __stcs(ap, __ldcs(bp));
    }
  }
}

}  // namespace

extern "C" int bricks_copy_array(bricks_ctx *ctx, int to_brick, const long dimlist[3], const long padding[3],
                                 const long ghost[3], double *arr_dev, const unsigned *grid_dev,
                                 double *bricks_dev, size_t step, int stream) {
  BRK_ARG(ctx && dimlist && padding && ghost && arr_dev && grid_dev && bricks_dev);
  BRK_ARG(stream >= 0 && stream < bricks::kNumStreams && step >= 512);
  ArrayCopyParams p;
  long sa = 1, sb = 1, strideA[3], strideB[3];
  for (int d = 0; d < 3; ++d) {
    BRK_ARG(dimlist[d] > 0 && dimlist[d] % 8 == 0 && ghost[d] >= 0 && ghost[d] % 8 == 0 && padding[d] >= 0);
    strideA[d] = sa;
    strideB[d] = sb;
    sa *= dimlist[d] + 2 * (padding[d] + ghost[d]);
    sb *= (dimlist[d] + 2 * ghost[d]) / 8;
    /* iter_grid visits cells [ghost/tile, (dimlist+ghost)/tile): the ghost band is skipped */
    p.lo[d] = (unsigned)(ghost[d] / 8);
    p.n[d] = (unsigned)(dimlist[d] / 8);
    p.off[d] = padding[d];
  }
  BRK_ARG(p.n[1] < 65536 && p.n[2] < 65536);
  p.arr = arr_dev;
  p.bricks = bricks_dev;
  p.grid = grid_dev;
  p.sa1 = strideA[1];
  p.sa2 = strideA[2];
  p.sb1 = strideB[1];
  p.sb2 = strideB[2];
  p.step = step;
  BRK_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t s = ctx->streams[stream];
  const dim3 grid((p.n[0] + kStrip - 1) / kStrip, p.n[1], p.n[2]);
  const bool vec = ((uintptr_t)arr_dev % 16 == 0) && ((uintptr_t)bricks_dev % 16 == 0) && padding[0] % 2 == 0 &&
                   strideA[1] % 2 == 0 && strideA[2] % 2 == 0 && step % 2 == 0;
  if (vec) {
    if (to_brick) array_copy_kernel<2, true><<<grid, kThreads, 0, s>>>(p);
    else array_copy_kernel<2, false><<<grid, kThreads, 0, s>>>(p);
  } else {
    if (to_brick) array_copy_kernel<1, true><<<grid, kThreads, 0, s>>>(p);
    else array_copy_kernel<1, false><<<grid, kThreads, 0, s>>>(p);
  }
  BRK_CUDA(cudaGetLastError());
  ++bricks::g_launches;
  return 0;
}
