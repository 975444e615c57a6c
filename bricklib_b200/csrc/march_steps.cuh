// Per-stage arithmetic of the marching kernels.  MODE: 0 = pre-roll (R planes below the
// column: only their contribution to later output planes), 1 = main stage (NZ planes of a
// brick layer), 2 = post-roll (R planes above the column: only their contribution to earlier
// planes).  `carry` holds the 2R partially summed output planes that straddle stages:
// carry[a] <-> output plane (a - R) relative to the first input plane of the next stage.
//
// Stencil definitions followed: stencils/7pt.py:19-26 (VAR=1: coeff[0..6] = centre, i+1, i-1,
// j+1, j-1, k+1, k-1), stencils/mpi7pt.py:17-24, mpi13pt.py:18-31, mpi25pt.py:20-45 (VAR=0:
// c[r] for |offset| = r along one axis), stencils/mpi125pt.py:12-161 (10 coefficients indexed by
// the sorted |offset| triple).  Summation order differs from the generated reference code, so
// results agree to rounding (<= 1e-12 relative is the bar), not bit for bit.
#pragma once
#include "march_dev.cuh"

namespace dev {

template <typename T = double>
struct OutPtrs {
  T *cur, *prev; /* lane's quad in plane 0 of the current / previous layer's brick */
  bool curv, prevv;
};

/* store a finished output plane, branch-free (see st_quad_if) */
template <int R, int MODE, typename T>
__device__ __forceinline__ void store_done_nb(const OutPtrs<T> &o, int zbase, int z, const T (&v)[4]) {
  if (MODE == 0) return;
  const int op = zbase + z - R;
  const bool here = op >= 0;
  st_quad_if((here ? o.cur : o.prev) + (here ? op : op + 8) * 64, v, here ? o.curv : o.prevv);
}

template <int R, int MODE, typename T>
__device__ __forceinline__ void store_done(const OutPtrs<T> &o, int zbase, int z, const T (&v)[4]) {
  store_done_nb<R, MODE, T>(o, zbase, z, v);
}

/* i-neighbours of the lane's quad: R values on each side, by shuffle from the adjacent lane or,
 * on the footprint edge, from the i-face piece. */
template <int R, int HW, typename T>
__device__ __forceinline__ void gather_i(const T (&x)[4], const T (&h)[HW], bool halo, int bb,
                                         int srcL, int srcR, T (&e)[4 + 2 * R]) {
#pragma unroll
  for (int m = 0; m < 4; ++m) e[R + m] = x[m];
#pragma unroll
  for (int r = 0; r < R; ++r) {
    const T l = __shfl_sync(0xffffffffu, x[4 - R + r], srcL);
    const T g = __shfl_sync(0xffffffffu, x[r], srcR);
    e[r] = (halo && bb == 0) ? h[HW - R + r] : l;
    e[R + 4 + r] = (halo && bb == 1) ? h[r] : g;
  }
}

/* Every lane loads the i-face row of its (brick, row): the two lanes of a row read the same 16
 * bytes (a broadcast, no extra wavefront), so the load needs neither a predicate nor a zero fill;
 * only the lane on the footprint edge uses the value (gather_i). */
template <int HW>
__device__ __forceinline__ void ld_halo(const char *st, int off, bool /*halo*/, double (&h)[HW]) {
#pragma unroll
  for (int q = 0; q < HW / 2; ++q) {
    const double2 v = *reinterpret_cast<const double2 *>(st + off + 16 * q);
    h[2 * q] = v.x;
    h[2 * q + 1] = v.y;
  }
}
template <int HW>
__device__ __forceinline__ void ld_halo(const char *st, int off, bool /*halo*/, float (&h)[HW]) {
  static_assert(HW == 4, "fp32 i-face rows are four elements = one 16-byte chunk");
  const float4 v = *reinterpret_cast<const float4 *>(st + off);
  h[0] = v.x, h[1] = v.y, h[2] = v.z, h[3] = v.w;
}

template <int N, typename T>
__device__ __forceinline__ void relu(T (&v)[N]) {
#pragma unroll
  for (int m = 0; m < N; ++m) v[m] = fmax(v[m], (T)0.0);
}

/* ------------------------------------------------------------------ star stencils
 * VAR 0: one coefficient per distance (mpi*pt.py); 1: 7pt.py's seven coefficients; 2: cond.py, the
 * same seven on max(x, 0) inputs with |.| on the result (stencils/cond.py:23-30). */
template <int R, int VAR, int P, int NZ, int MODE, typename T>
__device__ __forceinline__ void star_step(const char *st, const LaneK<R, false, P, T> &lk,
                                          const T *__restrict__ c, T (&carry)[2 * R][4],
                                          const OutPtrs<T> &o, int zbase) {
  constexpr int HW = StageLayout<R, false, P, T>::HW;
  T acc[NZ + 2 * R][4];
#pragma unroll
  for (int a = 0; a < NZ + 2 * R; ++a)
#pragma unroll
    for (int m = 0; m < 4; ++m) acc[a][m] = a < 2 * R ? carry[a][m] : 0.0;

#pragma unroll
  for (int z = 0; z < NZ; ++z) {
    T x[4];
    ld_quad(st, lk.row0[R] + z * lk.rpitch[R], lk.row1[R] + z * lk.rpitch[R], x);
    if (VAR == 2) relu(x); /* cond.py: every input through max(x, 0) */
    if (MODE == 1) {
      T h[HW], e[4 + 2 * R];
      ld_halo<HW>(st, lk.hrow[R] + z * lk.hpitch[R], lk.halo, h);
      if (VAR == 2) relu(h);
      gather_i<R, HW, T>(x, h, lk.halo, lk.bb, lk.srcL, lk.srcR, e);
      T s[4];
#pragma unroll
      for (int m = 0; m < 4; ++m) s[m] = c[0] * x[m];
#pragma unroll
      for (int r = 1; r <= R; ++r) {
        T up[4], dn[4];
        if (StageLayout<R, false, P, T>::JHX && (z & 1))
          ld_quad(st, (lk.row0[R + r] + z * lk.rpitch[R + r]) ^ lk.jx[R + r],
                  (lk.row1[R + r] + z * lk.rpitch[R + r]) ^ lk.jx[R + r], up);
        else
          ld_quad(st, lk.row0[R + r] + z * lk.rpitch[R + r], lk.row1[R + r] + z * lk.rpitch[R + r], up);
        if (StageLayout<R, false, P, T>::JLX && (z & 1)) /* swizzled -j face rows, odd planes (march_dev.cuh) */
          ld_quad(st, (lk.row0[R - r] + z * lk.rpitch[R - r]) ^ lk.jx[R - r],
                  (lk.row1[R - r] + z * lk.rpitch[R - r]) ^ lk.jx[R - r], dn);
        else
          ld_quad(st, lk.row0[R - r] + z * lk.rpitch[R - r], lk.row1[R - r] + z * lk.rpitch[R - r], dn);
        if (VAR == 2) relu(up), relu(dn);
#pragma unroll
        for (int m = 0; m < 4; ++m) {
          if (VAR) {
            s[m] += c[1] * e[R + m + 1];
            s[m] += c[2] * e[R + m - 1];
            s[m] += c[3] * up[m];
            s[m] += c[4] * dn[m];
          } else {
            s[m] += c[r] * ((e[R + m + r] + e[R + m - r]) + (up[m] + dn[m]));
          }
        }
      }
#pragma unroll
      for (int m = 0; m < 4; ++m) acc[z + R][m] += s[m];
    }
#pragma unroll
    for (int r = 1; r <= R; ++r)
#pragma unroll
      for (int m = 0; m < 4; ++m) {
        /* input plane z feeds output planes z-r (needs x(k+r)) and z+r (needs x(k-r)).  Explicit
         * FMAs: left to the compiler, the product shared by both directions becomes one rounded
         * DMUL in the main stage but an FMA in the rolls, and results would depend (in the last
         * bit) on where a column starts. */
        if (MODE != 0) acc[z + R - r][m] = fma(VAR ? c[5] : c[r], x[m], acc[z + R - r][m]);
        if (MODE != 2) {
          if (r == R)
            acc[z + R + r][m] = (VAR ? c[6] : c[r]) * x[m];
          else
            acc[z + R + r][m] = fma(VAR ? c[6] : c[r], x[m], acc[z + R + r][m]);
        }
      }
    if (VAR == 2) { /* cond.py: If(calc > 0, calc, -calc) */
      T v[4];
#pragma unroll
      for (int m = 0; m < 4; ++m) v[m] = fabs(acc[z][m]);
      store_done<R, MODE, T>(o, zbase, z, v);
    } else {
      store_done<R, MODE, T>(o, zbase, z, acc[z]);
    }
  }
#pragma unroll
  for (int a = 0; a < 2 * R; ++a)
#pragma unroll
    for (int m = 0; m < 4; ++m) carry[a][m] = acc[NZ + a][m];
}

/* ------------------------------------------------------------------ 27-point cube, 27 coefficients
 * d3pt27 (stencils/3axis.cpp:165-247).  w[(dk+1)*9 + (dj+1)*3 + (di+1)] is the weight of
 * in(k+dk, j+dj, i+di); an input plane z feeds output plane z-1 with the dk = +1 block, z with the
 * dk = 0 block and z+1 with the dk = -1 block: 27 FMAs per point. */
template <int P, int NZ, int MODE, typename T>
__device__ __forceinline__ void cube27_step(const char *st, const LaneK<1, true, P, T> &lk,
                                            const T *__restrict__ w, T (&carry)[2][4],
                                            const OutPtrs<T> &o, int zbase) {
  constexpr int R = 1, HW = StageLayout<1, true, P, T>::HW;
  T acc[NZ + 2][4];
#pragma unroll
  for (int a = 0; a < NZ + 2; ++a)
#pragma unroll
    for (int m = 0; m < 4; ++m) acc[a][m] = a < 2 ? carry[a][m] : 0.0;

#pragma unroll
  for (int z = 0; z < NZ; ++z) {
    T e[3][6]; /* rows j-1, j, j+1; columns i-1 .. i+4 of the lane's quad */
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      T x[4], h[HW];
      ld_quad(st, lk.row0[a] + z * lk.rpitch[a], lk.row1[a] + z * lk.rpitch[a], x);
      ld_halo<HW>(st, lk.hrow[a] + z * lk.hpitch[a], lk.halo, h);
      gather_i<R, HW, T>(x, h, lk.halo, lk.bb, lk.srcL, lk.srcR, e[a]);
    }
#pragma unroll
    for (int m = 0; m < 4; ++m) {
#pragma unroll
      for (int blk = 0; blk < 3; ++blk) {
        /* block blk = dk + 1 goes to output plane z - dk, i.e. acc[z + 2 - blk] */
        if ((MODE == 0 && blk != 0) || (MODE == 2 && blk != 2)) continue;
        T s = blk == 0 ? 0.0 : acc[z + 2 - blk][m];
#pragma unroll
        for (int dj = 0; dj < 3; ++dj)
#pragma unroll
          for (int di = 0; di < 3; ++di) s = fma(w[blk * 9 + dj * 3 + di], e[dj][m + di], s);
        acc[z + 2 - blk][m] = s;
      }
    }
    store_done<R, MODE, T>(o, zbase, z, acc[z]);
  }
#pragma unroll
  for (int a = 0; a < 2; ++a)
#pragma unroll
    for (int m = 0; m < 4; ++m) carry[a][m] = acc[NZ + a][m];
}

/* One input plane of the cube: e[b][0..7] = v_b over the lane's quad and two columns either side.
 * Adds the plane's contributions to the five output planes it feeds (acc[z..z+4]). */
template <int NZ, int MODE, typename T>
__device__ __forceinline__ void cube_scatter(const T (&e)[3][8], const T *__restrict__ c,
                                             T (&acc)[NZ + 4][4], int z) {
#pragma unroll
    for (int m = 0; m < 4; ++m) {
      T t[3][3]; /* t[a][b]: i-offset a, j-offset b */
#pragma unroll
      for (int b = 0; b < 3; ++b) {
        t[0][b] = e[b][2 + m];
        t[1][b] = e[b][3 + m] + e[b][1 + m];
        t[2][b] = e[b][4 + m] + e[b][m];
      }
      const T q00 = t[0][0], q11 = t[1][1], q22 = t[2][2];
      const T q01 = t[0][1] + t[1][0], q02 = t[0][2] + t[2][0], q12 = t[1][2] + t[2][1];
      /* coefficient classes of stencils/mpi125pt.py:12-33: c[0]=(0,0,0) c[1]=(0,0,1) c[2]=(0,0,2)
       * c[3]=(0,1,1) c[4]=(0,1,2) c[5]=(0,2,2) c[6]=(1,1,1) c[7]=(1,1,2) c[8]=(1,2,2) c[9]=(2,2,2) */
      T g[3];
      if (MODE == 1) { /* the in-plane part goes straight into its accumulator: six FMAs */
        T a0 = acc[z + 2][m];
        a0 = fma(c[0], q00, a0);
        a0 = fma(c[1], q01, a0);
        a0 = fma(c[2], q02, a0);
        a0 = fma(c[3], q11, a0);
        a0 = fma(c[4], q12, a0);
        acc[z + 2][m] = fma(c[5], q22, a0);
      }
      g[1] = c[1] * q00 + c[3] * q01 + c[4] * q02 + c[6] * q11 + c[7] * q12 + c[8] * q22;
      g[2] = c[2] * q00 + c[4] * q01 + c[5] * q02 + c[7] * q11 + c[8] * q12 + c[9] * q22;
      if (MODE != 0) {
        acc[z + 1][m] += g[1];
        acc[z][m] += g[2];
      }
      if (MODE != 2) {
        acc[z + 3][m] += g[1];
        acc[z + 4][m] = g[2];
      }
    }
}

/* ------------------------------------------------------------------ 125-point cube
 * Factorisation: per input plane, with v_b = x(j+b)+x(j-b) (v_0 = x),
 * t_ab = v_b(i+a)+v_b(i-a), Q_ab = t_ab + t_ba (a<b), the plane's contribution to output plane
 * z+-c is G_c = sum_{a<=b} C[class(a,b,c)] * Q_ab.  ~35 fp64 ops per point instead of 125. */
template <int P, int NZ, int MODE, typename T>
__device__ __forceinline__ void cube_step(const char *st, const LaneK<2, true, P, T> &lk,
                                          const T *__restrict__ c, T (&carry)[4][4],
                                          const OutPtrs<T> &o, int zbase) {
  constexpr int R = 2, HW = StageLayout<2, true, P, T>::HW;
#ifdef MARCH_EXP_NOMATH /* experiment build only: the load/store skeleton without arithmetic */
  if (MODE == 1) {
#pragma unroll
    for (int z = 0; z < NZ; ++z) {
      T x[4];
      ld_quad(st, lk.row0[R] + z * lk.rpitch[R], lk.row1[R] + z * lk.rpitch[R], x);
      if (o.curv) st_quad(o.cur + (zbase + z) * 64, x);
    }
  }
  return;
#endif
  T acc[NZ + 4][4];
#pragma unroll
  for (int a = 0; a < NZ + 4; ++a)
#pragma unroll
    for (int m = 0; m < 4; ++m) acc[a][m] = a < 4 ? carry[a][m] : 0.0;

#pragma unroll
  for (int z = 0; z < NZ; ++z) {
    /* v_b on the lane's quad and on its i-face pair */
    T v[3][4], hv[3][HW];
    {
      T xr[5][4], hr[5][HW];
#pragma unroll
      for (int a = 0; a < 5; ++a) {
        ld_quad(st, lk.row0[a] + z * lk.rpitch[a], lk.row1[a] + z * lk.rpitch[a], xr[a]);
        ld_halo<HW>(st, lk.hrow[a] + z * lk.hpitch[a], lk.halo, hr[a]);
      }
#pragma unroll
      for (int m = 0; m < 4; ++m) {
        v[0][m] = xr[2][m];
        v[1][m] = xr[1][m] + xr[3][m];
        v[2][m] = xr[0][m] + xr[4][m];
      }
#pragma unroll
      for (int q = 0; q < HW; ++q) {
        hv[0][q] = hr[2][q];
        hv[1][q] = hr[1][q] + hr[3][q];
        hv[2][q] = hr[0][q] + hr[4][q];
      }
    }
    T e[3][8];
#pragma unroll
    for (int b = 0; b < 3; ++b) gather_i<R, HW, T>(v[b], hv[b], lk.halo, lk.bb, lk.srcL, lk.srcR, e[b]);

    cube_scatter<NZ, MODE, T>(e, c, acc, z);
    store_done<R, MODE, T>(o, zbase, z, acc[z]);
  }
#pragma unroll
  for (int a = 0; a < 4; ++a)
#pragma unroll
    for (int m = 0; m < 4; ++m) carry[a][m] = acc[NZ + a][m];
}

}  // namespace dev
