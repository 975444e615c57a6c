// brickcompare.h -- brick-vs-array comparison.
//
// Drop-in for the reference's include/brickcompare.h:30-79 + include/cmpconst.h:9
// (compareBrick: pass iff every visited element differs by < BRICK_TOLERANCE absolutely or
// relative to |a|+|b|), plus the strict measure this project is judged on: the maximum
// relative error (bar: <= 1e-12 in fp64).
#ifndef BRICKS_B200_BRICKCOMPARE_H
#define BRICKS_B200_BRICKCOMPARE_H

#include <cmath>
#include <vector>

#include "bricksetup.h"

#ifndef BRICK_TOLERANCE
#define BRICK_TOLERANCE 1e-6
#endif

/// max over visited elements of |brick - array| / max(|array|, tiny)
template <unsigned dims, typename T>
inline double maxRelErrBrick(const std::vector<long> &dimlist, const std::vector<long> &padding,
                             const std::vector<long> &ghost, bElem *arr, unsigned *grid_ptr, T &brick) {
  double worst = 0.0;
  // iter_grid parallelises over the outer brick axis; the rare improving element takes a lock
  iter_grid<dims>(dimlist, padding, ghost, arr, grid_ptr, brick, [&worst](bElem &e, bElem *a) {
    const double ref = std::abs(*a);
    const double r = std::abs(e - *a) / (ref > 1e-300 ? ref : 1e-300);
    if (r > worst || r != r) {
#pragma omp critical(bricks_maxrel)
      if (r > worst || r != r) worst = (r != r) ? INFINITY : r;
    }
  });
  return worst;
}

template <unsigned dims, typename T>
inline bool compareBrick(const std::vector<long> &dimlist, const std::vector<long> &padding,
                         const std::vector<long> &ghost, bElem *arr, unsigned *grid_ptr, T &brick) {
  int bad = 0;
  iter_grid<dims>(dimlist, padding, ghost, arr, grid_ptr, brick, [&bad](bElem &e, bElem *a) {
    const double diff = std::abs(e - *a);
    const bool ok = (diff < BRICK_TOLERANCE) || (diff < (std::abs(e) + std::abs(*a)) * BRICK_TOLERANCE);
    if (!ok) {
#pragma omp atomic write
      bad = 1;
    }
  });
  return bad == 0;
}

template <unsigned dims, typename T>
inline bool compareBrick(const std::vector<long> &dimlist, bElem *arr, unsigned *grid_ptr, T &brick) {
  std::vector<long> zero(dimlist.size(), 0);
  return compareBrick<dims, T>(dimlist, zero, zero, arr, grid_ptr, brick);
}

#endif  // BRICKS_B200_BRICKCOMPARE_H
