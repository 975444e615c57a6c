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

// ---------------------------------------------------------------------------------------------
//   bricks_array_stencil7   the array-layout baseline of the reference's single-GPU driver:
//                           d3pt7_arr / d3pt7_arr_warp (stencils/3axis.cu:39-60), the 7-point stencil
//                           with seven coefficients on a conventional padded array, for an on-box
//                           brick-vs-array comparison.  NOT the product path; it exists so that the
//                           "Arr" line of d3pt7cu has a B200 counterpart.
// A block owns a 32 (i) x 8 (j) column and marches it along k: the k-1 / k / k+1 values of a thread's
// point stay in registers, so every element is fetched from global memory once as a centre and is an
// L1 hit for its four in-plane neighbours.
namespace {

struct Array7Params {
  const double *in;
  double *out;
  long s1, s2;          /* strides of j and k in elements */
  long lo[3], n[3];     /* box origin and extent, i first */
  double c[7];
  int kchunk;
};

__global__ void __launch_bounds__(256) array7_kernel(const Array7Params p) {
  const long i = (long)blockIdx.x * 32 + threadIdx.x, j = (long)blockIdx.y * 8 + threadIdx.y;
  if (i >= p.n[0] || j >= p.n[1]) return;
  const long k0 = (long)blockIdx.z * p.kchunk;
  const long k1 = k0 + p.kchunk < p.n[2] ? k0 + p.kchunk : p.n[2];
  const double *a = p.in + (p.lo[2] + k0) * p.s2 + (p.lo[1] + j) * p.s1 + p.lo[0] + i;
  double *o = p.out + (p.lo[2] + k0) * p.s2 + (p.lo[1] + j) * p.s1 + p.lo[0] + i;
  double below = a[-p.s2], here = a[0];
  for (long k = k0; k < k1; ++k) {
    const double above = a[p.s2];
    /* term order of the reference: k+1, k-1, j+1, j-1, i+1, i-1, centre (3axis.cu:44-47) */
    const double r = p.c[5] * above + p.c[6] * below + p.c[3] * a[p.s1] + p.c[4] * a[-p.s1] + p.c[1] * a[1] +
                     p.c[2] * a[-1] + p.c[0] * here;
    __stcs(o, r);
    below = here;
    here = above;
    a += p.s2;
    o += p.s2;
  }
}

}  // namespace

extern "C" int bricks_array_stencil7(bricks_ctx *ctx, const double *in_dev, double *out_dev, const long ext[3],
                                     const long lo[3], const long hi[3], const double *coeff_host, int stream) {
  BRK_ARG(ctx && in_dev && out_dev && ext && lo && hi && coeff_host && in_dev != out_dev);
  BRK_ARG(stream >= 0 && stream < bricks::kNumStreams);
  Array7Params p;
  for (int d = 0; d < 3; ++d) {
    BRK_ARG(lo[d] >= 1 && hi[d] > lo[d] && hi[d] <= ext[d] - 1); /* one element of halo inside the array */
    p.lo[d] = lo[d];
    p.n[d] = hi[d] - lo[d];
  }
  p.in = in_dev;
  p.out = out_dev;
  p.s1 = ext[0];
  p.s2 = ext[0] * ext[1];
  for (int t = 0; t < 7; ++t) p.c[t] = coeff_host[t];
  p.kchunk = 32;
  BRK_ARG((p.n[1] + 7) / 8 < 65536 && (p.n[2] + p.kchunk - 1) / p.kchunk < 65536);
  BRK_CUDA(cudaSetDevice(ctx->device));
  const dim3 grid((unsigned)((p.n[0] + 31) / 32), (unsigned)((p.n[1] + 7) / 8), (unsigned)((p.n[2] + p.kchunk - 1) / p.kchunk));
  array7_kernel<<<grid, dim3(32, 8), 0, ctx->streams[stream]>>>(p);
  BRK_CUDA(cudaGetLastError());
  ++bricks::g_launches;
  return 0;
}
