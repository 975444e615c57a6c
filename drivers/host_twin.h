// host_twin.h -- the ARRAY twin the drivers validate against: the stencil applied to a conventional
// padded array on the host, with a periodic halo exchange between the ranks' arrays.
//
// The reference's MPI drivers validate the brick result against the array path (weak/main.cu:143-214
// builds it with exchangeArr + ST_GPU, :321-323 compares with compareBrick; strong/main.cpp:402-476
// does the same over all subdomains).  Ranks are host threads of one process here, so "exchangeArr"
// (include/array-mpi.h) is a copy between the ranks' arrays after a barrier.  Nothing in this file
// touches a device or the brick layout: it is an independent statement of stencils/fake.h:41-352
// (ST_CPU) and of the periodic process grid.
#ifndef BRICKS_B200_DRIVERS_HOST_TWIN_H
#define BRICKS_B200_DRIVERS_HOST_TWIN_H

#include <algorithm>
#include <cstdlib>
#include <vector>

namespace twin {

struct Tap {
  int dk, dj, di;
  double c;
};

/// point set and coefficients of mpi7pt.py / mpi13pt.py / mpi25pt.py / mpi125pt.py with the constants of
/// stencils/fake.h:11-33
inline std::vector<Tap> taps(int which) {
  std::vector<Tap> t;
  if (which == 125) {
    // MPI_C0..C9, indexed by the sorted |offset| triple: (0,0,0) (0,0,1) (0,0,2) (0,1,1) (0,1,2) (0,2,2)
    // (1,1,1) (1,1,2) (1,2,2) (2,2,2)  (stencils/mpi125pt.py:12-33)
    const double C[10] = {0.1, 0.04, 0.03, 0.01, 0.006, 0.004, 0.005, 0.002, 0.003, 0.001};
    const int cls[3][3][3] = {{{0, 1, 2}, {1, 3, 4}, {2, 4, 5}}, {{1, 3, 4}, {3, 6, 7}, {4, 7, 8}}, {{2, 4, 5}, {4, 7, 8}, {5, 8, 9}}};
    for (int k = -2; k <= 2; ++k)
      for (int j = -2; j <= 2; ++j)
        for (int i = -2; i <= 2; ++i) t.push_back(Tap{k, j, i, C[cls[std::abs(k)][std::abs(j)][std::abs(i)]]});
    return t;
  }
  const double c7[2] = {0.4, 0.1}, c13[3] = {0.4, 0.07, 0.03}, c25[5] = {0.1, 0.06, 0.045, 0.03, 0.015};
  const double *c = which == 7 ? c7 : (which == 13 ? c13 : c25);
  const int R = which == 7 ? 1 : (which == 13 ? 2 : 4);
  t.push_back(Tap{0, 0, 0, c[0]});
  for (int r = 1; r <= R; ++r) {
    t.push_back(Tap{0, 0, r, c[r]}), t.push_back(Tap{0, 0, -r, c[r]});
    t.push_back(Tap{0, r, 0, c[r]}), t.push_back(Tap{0, -r, 0, c[r]});
    t.push_back(Tap{r, 0, 0, c[r]}), t.push_back(Tap{-r, 0, 0, c[r]});
  }
  return t;
}

/// out = stencil(in) on [lo, hi)^3 of an array with `ext` elements per axis (i fastest)
inline void sweep(int which, const double *in, double *out, const long ext[3], const long lo[3], const long hi[3]) {
  const std::vector<Tap> t = taps(which);
  const long s1 = ext[0], s2 = ext[0] * ext[1];
#pragma omp parallel for collapse(2)
  for (long k = lo[2]; k < hi[2]; ++k)
    for (long j = lo[1]; j < hi[1]; ++j)
      for (long i = lo[0]; i < hi[0]; ++i) {
        double s = 0.0;
        for (const Tap &p : t) s += p.c * in[(k + p.dk) * s2 + (j + p.dj) * s1 + i + p.di];
        out[k * s2 + j * s1 + i] = s;
      }
}

/// One rank's array with a ghost band of `gz` elements: ext = dom + 2 * gz.  fill_ghost copies every
/// ghost element from the array of the rank that owns it on the periodic process grid;
/// owner(sk, sj, si) names that rank for a displacement of -1 / 0 / +1 subdomains per axis.
template <typename OwnerFn>
inline void fill_ghost(double *mine, const std::vector<double *> &all, const long dom[3], long gz, OwnerFn owner) {
  const long ext[3] = {dom[0] + 2 * gz, dom[1] + 2 * gz, dom[2] + 2 * gz};
  const long s1 = ext[0], s2 = ext[0] * ext[1];
#pragma omp parallel for
  for (long k = 0; k < ext[2]; ++k) {
    const int sk = k < gz ? -1 : (k >= gz + dom[2] ? 1 : 0);
    for (long j = 0; j < ext[1]; ++j) {
      const int sj = j < gz ? -1 : (j >= gz + dom[1] ? 1 : 0);
      for (long i = 0; i < ext[0]; ++i) {
        const int si = i < gz ? -1 : (i >= gz + dom[0] ? 1 : 0);
        if (!sk && !sj && !si) {
          i = gz + dom[0] - 1;  // skip the interior of this row
          continue;
        }
        const double *src = all[owner(sk, sj, si)];
        mine[k * s2 + j * s1 + i] = src[(k - sk * dom[2]) * s2 + (j - sj * dom[1]) * s1 + (i - si * dom[0])];
      }
    }
  }
}

/// The 4-D twin (weak/main-4d.cpp:171-193 array_stencil with exchangeArr<4>): out = 0.2 * in + 0.1 * (the
/// eight axis neighbours) on [lo, hi)^4 of an array with `ext` elements per axis (i fastest)
inline void sweep4d(const double *in, double *out, const long ext[4], const long lo[4], const long hi[4]) {
  const long s1 = ext[0], s2 = s1 * ext[1], s3 = s2 * ext[2];
#pragma omp parallel for collapse(2)
  for (long l = lo[3]; l < hi[3]; ++l)
    for (long k = lo[2]; k < hi[2]; ++k)
      for (long j = lo[1]; j < hi[1]; ++j)
        for (long i = lo[0]; i < hi[0]; ++i) {
          const long p = l * s3 + k * s2 + j * s1 + i;
          out[p] = 0.2 * in[p] + 0.1 * (in[p + 1] + in[p - 1] + in[p + s1] + in[p - s1] + in[p + s2] + in[p - s2] +
                                        in[p + s3] + in[p - s3]);
        }
}

/// fill_ghost for four axes; owner(sl, sk, sj, si)
template <typename OwnerFn>
inline void fill_ghost4d(double *mine, const std::vector<double *> &all, const long dom[4], long gz, OwnerFn owner) {
  const long ext[4] = {dom[0] + 2 * gz, dom[1] + 2 * gz, dom[2] + 2 * gz, dom[3] + 2 * gz};
  const long s1 = ext[0], s2 = s1 * ext[1], s3 = s2 * ext[2];
  auto side = [&](long x, int a) { return x < gz ? -1 : (x >= gz + dom[a] ? 1 : 0); };
#pragma omp parallel for collapse(2)
  for (long l = 0; l < ext[3]; ++l)
    for (long k = 0; k < ext[2]; ++k) {
      const int sl = side(l, 3), sk = side(k, 2);
      for (long j = 0; j < ext[1]; ++j) {
        const int sj = side(j, 1);
        for (long i = 0; i < ext[0]; ++i) {
          const int si = side(i, 0);
          if (!sl && !sk && !sj && !si) {
            i = gz + dom[0] - 1;  // skip the interior of this row
            continue;
          }
          const double *src = all[owner(sl, sk, sj, si)];
          mine[l * s3 + k * s2 + j * s1 + i] =
              src[(l - sl * dom[3]) * s3 + (k - sk * dom[2]) * s2 + (j - sj * dom[1]) * s1 + (i - si * dom[0])];
        }
      }
    }
}

}  // namespace twin

#endif  // BRICKS_B200_DRIVERS_HOST_TWIN_H
