// bricksetup.h -- single-node brick grid setup and array <-> brick copies.
//
// Drop-in for the reference's include/bricksetup.h: init_grid<dims> (:73-90, identity
// numbering, adjacency from +-stride offsets, out-of-range -> brick 0), iter_grid (:139-160),
// copyToBrick / copyFromBrick (:172-221).  Same names and argument meaning (dimension lists
// are contiguous-dimension-first); implemented with flat odometer loops instead of
// tag-dispatched recursion.
#ifndef BRICKS_B200_BRICKSETUP_H
#define BRICKS_B200_BRICKSETUP_H

#include <algorithm>
#include <vector>

#include "brick.h"

/// Number the bricks of a dimlist[0] x dimlist[1] x ... grid 0,1,2,.. (axis 0 fastest) and
/// build the adjacency list.  A neighbour whose FLAT index falls outside [0, size) maps to
/// brick 0; neighbours that merely wrap around a row keep their flat index (this is the
/// reference's behaviour, bricksetup.h:44-47: only the pointer range is checked).
template <unsigned dims>
BrickInfo<dims> init_grid(unsigned *&grid_ptr, const std::vector<long> &dimlist) {
  long size = 1;
  std::vector<long> stride(dims);
  for (unsigned a = 0; a < dims; ++a) {
    stride[a] = size;
    size *= dimlist[a];
  }
  grid_ptr = (unsigned *)malloc(size * sizeof(unsigned));
  for (long p = 0; p < size; ++p) grid_ptr[p] = (unsigned)p;
  BrickInfo<dims> info((unsigned)size);
  constexpr unsigned NN = static_power<3, dims>::value;
  // flat offset of every neighbour slot; slot digits are outermost axis first
  long noff[NN];
  for (unsigned n = 0; n < NN; ++n) {
    long off = 0;
    unsigned rem = n;
    for (unsigned a = 0; a < dims; ++a) {  // least significant digit = axis 0
      off += ((long)(rem % 3) - 1) * stride[a];
      rem /= 3;
    }
    noff[n] = off;
  }
#pragma omp parallel for
  for (long p = 0; p < size; ++p)
    for (unsigned n = 0; n < NN; ++n) {
      const long q = p + noff[n];
      info.adj[p][n] = (q >= 0 && q < size) ? grid_ptr[q] : 0;
    }
  return info;
}

/// Visit every element of the bricks covering the region `dimlist` (+ `ghost` on each side)
/// of an array that carries `padding + ghost` extra elements on each side, calling
/// f(brick_element_ref, array_element_ptr).  grid_ptr is the brick-index grid over
/// (dimlist + 2*ghost) / tile cells.  Bricks inside the ghost band are NOT visited (the
/// reference iterates s in [ghost/tile, (dimlist+ghost)/tile), bricksetup.h:108-124).
template <unsigned dims, typename F, typename T, unsigned... BDims>
inline void iter_grid(const std::vector<long> &dimlist, const std::vector<long> &padding,
                      const std::vector<long> &ghost, bElem *arr, unsigned *grid_ptr,
                      Brick<Dim<BDims...>, T> &brick, F f) {
  static_assert(dims == 3, "iter_grid: this build implements the 3-D case");
  std::vector<long> tile = {BDims...};
  std::reverse(tile.begin(), tile.end());  // contiguous axis first
  long strideA[3], strideB[3], sA = 1, sB = 1;
  for (unsigned a = 0; a < 3; ++a) {
    strideA[a] = sA;
    strideB[a] = sB;
    sA *= dimlist[a] + 2 * (padding[a] + ghost[a]);
    sB *= (dimlist[a] + 2 * ghost[a]) / tile[a];
  }
  long lo[3], hi[3];
  for (unsigned a = 0; a < 3; ++a) {
    lo[a] = ghost[a] / tile[a];
    hi[a] = (dimlist[a] + ghost[a]) / tile[a];
  }
#pragma omp parallel for
  for (long bk = lo[2]; bk < hi[2]; ++bk)
    for (long bj = lo[1]; bj < hi[1]; ++bj)
      for (long bi = lo[0]; bi < hi[0]; ++bi) {
        const unsigned b = grid_ptr[bi * strideB[0] + bj * strideB[1] + bk * strideB[2]];
        bElem *base = arr + (padding[0] + bi * tile[0]) * strideA[0] + (padding[1] + bj * tile[1]) * strideA[1] +
                      (padding[2] + bk * tile[2]) * strideA[2];
        for (long k = 0; k < tile[2]; ++k)
          for (long j = 0; j < tile[1]; ++j)
            for (long i = 0; i < tile[0]; ++i)
              f(brick[b][(unsigned)k][(unsigned)j][(unsigned)i], base + i * strideA[0] + j * strideA[1] + k * strideA[2]);
      }
}

template <unsigned dims, typename T>
inline void copyToBrick(const std::vector<long> &dimlist, const std::vector<long> &padding,
                        const std::vector<long> &ghost, bElem *arr, unsigned *grid_ptr, T &brick) {
  iter_grid<dims>(dimlist, padding, ghost, arr, grid_ptr, brick, [](bElem &e, bElem *a) { e = *a; });
}

template <unsigned dims, typename T>
inline void copyToBrick(const std::vector<long> &dimlist, bElem *arr, unsigned *grid_ptr, T &brick) {
  std::vector<long> zero(dimlist.size(), 0);
  copyToBrick<dims>(dimlist, zero, zero, arr, grid_ptr, brick);
}

template <unsigned dims, typename T>
inline void copyFromBrick(const std::vector<long> &dimlist, const std::vector<long> &padding,
                          const std::vector<long> &ghost, bElem *arr, unsigned *grid_ptr, T &brick) {
  iter_grid<dims>(dimlist, padding, ghost, arr, grid_ptr, brick, [](bElem &e, bElem *a) { *a = e; });
}

#endif  // BRICKS_B200_BRICKSETUP_H
