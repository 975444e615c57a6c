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

namespace bricks_detail {
/// brick[b][x_{n-1}]..[x_0] for a coordinate list given contiguous axis first
template <typename B>
inline bElem &brick_element(B &brick, unsigned b, const long *x) {
  static_assert(B::NDIMS >= 1 && B::NDIMS <= 4, "brick_element: 1 to 4 axes");
  if constexpr (B::NDIMS == 1)
    return brick[b][(unsigned)x[0]];
  else if constexpr (B::NDIMS == 2)
    return brick[b][(unsigned)x[1]][(unsigned)x[0]];
  else if constexpr (B::NDIMS == 3)
    return brick[b][(unsigned)x[2]][(unsigned)x[1]][(unsigned)x[0]];
  else
    return brick[b][(unsigned)x[3]][(unsigned)x[2]][(unsigned)x[1]][(unsigned)x[0]];
}
}  // namespace bricks_detail

/// Visit every element of the bricks covering the region `dimlist` (+ `ghost` on each side)
/// of an array that carries `padding + ghost` extra elements on each side, calling
/// f(brick_element_ref, array_element_ptr).  grid_ptr is the brick-index grid over
/// (dimlist + 2*ghost) / tile cells.  Bricks inside the ghost band are NOT visited (the
/// reference iterates s in [ghost/tile, (dimlist+ghost)/tile), bricksetup.h:108-124).
template <unsigned dims, typename F, typename T, unsigned... BDims>
inline void iter_grid(const std::vector<long> &dimlist, const std::vector<long> &padding,
                      const std::vector<long> &ghost, bElem *arr, unsigned *grid_ptr,
                      Brick<Dim<BDims...>, T> &brick, F f) {
  static_assert(dims == sizeof...(BDims), "iter_grid: one brick axis per array axis");
  std::vector<long> tile = {BDims...};
  std::reverse(tile.begin(), tile.end());  // contiguous axis first
  long strideA[dims], strideB[dims], lo[dims], cnt[dims], sA = 1, sB = 1, nbricks = 1, nelem = 1;
  for (unsigned a = 0; a < dims; ++a) {
    strideA[a] = sA;
    strideB[a] = sB;
    sA *= dimlist[a] + 2 * (padding[a] + ghost[a]);
    sB *= (dimlist[a] + 2 * ghost[a]) / tile[a];
    lo[a] = ghost[a] / tile[a];
    cnt[a] = (dimlist[a] + ghost[a]) / tile[a] - lo[a];
    nbricks *= cnt[a];
    nelem *= tile[a];
  }
  // one flat loop over the visited bricks, a flat loop over the elements of each
#pragma omp parallel for
  for (long n = 0; n < nbricks; ++n) {
    long rem = n, cell = 0;
    bElem *base = arr;
    for (unsigned a = 0; a < dims; ++a) {
      const long c = lo[a] + rem % cnt[a];
      rem /= cnt[a];
      cell += c * strideB[a];
      base += (padding[a] + c * tile[a]) * strideA[a];
    }
    const unsigned b = grid_ptr[cell];
    long x[dims];
    for (unsigned a = 0; a < dims; ++a) x[a] = 0;
    for (long e = 0; e < nelem; ++e) {
      long off = 0;
      for (unsigned a = 0; a < dims; ++a) off += x[a] * strideA[a];
      f(bricks_detail::brick_element(brick, b, x), base + off);
      unsigned a = 0;
      while (a < dims && ++x[a] == tile[a]) x[a++] = 0;  // odometer, contiguous axis fastest
    }
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
