// multiarray.h -- padded n-D array helpers.
//
// Drop-in for the reference's include/multiarray.h + src/multiarray.cpp:13-63
// (uninitArray / randomArray / zeroArray / compareArray, dimension list contiguous first).
// Header-only here.  The reference seeds a thread-private mt19937_64 from random_device, so
// its fields differ from run to run; parity needs identical synthetic grids, so
// randomArray(list, seed) is a counter-based U[0,1) generator (splitmix64 of the element
// index): the same values on every machine and for any thread count.  randomArray(list)
// keeps the reference signature and uses seed 0.
#ifndef BRICKS_B200_MULTIARRAY_H
#define BRICKS_B200_MULTIARRAY_H

#include <cmath>
#include <cstdint>
#include <vector>

#include "brick.h"

#ifndef BRICK_TOLERANCE
#define BRICK_TOLERANCE 1e-6
#endif

namespace bricks_detail {
inline uint64_t splitmix64(uint64_t x) {
  x += 0x9E3779B97F4A7C15ull;
  uint64_t z = x;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
/// element `idx` of stream `seed`; identical to bricklib_b200/synth.py:uniform
inline double uniform01(uint64_t seed, uint64_t idx) {
  const uint64_t key = splitmix64(seed * 0xD1342543DE82EF95ull + 1ull);
  return (double)(splitmix64(idx ^ key) >> 11) * (1.0 / 9007199254740992.0);
}
}  // namespace bricks_detail

inline bElem *uninitArray(const std::vector<long> &list, long &size) {
  size = 1;
  for (long n : list) size *= n;
  size_t bytes = ((size_t)size * sizeof(bElem) + ALIGN - 1) / ALIGN * ALIGN;
  return (bElem *)aligned_alloc(ALIGN, bytes);
}

inline bElem *randomArray(const std::vector<long> &list, uint64_t seed) {
  long size;
  bElem *arr = uninitArray(list, size);
#pragma omp parallel for
  for (long l = 0; l < size; ++l) arr[l] = (bElem)bricks_detail::uniform01(seed, (uint64_t)l);
  return arr;
}

inline bElem *randomArray(const std::vector<long> &list) { return randomArray(list, 0); }

inline bElem *zeroArray(const std::vector<long> &list) {
  long size;
  bElem *arr = uninitArray(list, size);
#pragma omp parallel for
  for (long l = 0; l < size; ++l) arr[l] = 0.0;
  return arr;
}

inline bool compareArray(const std::vector<long> &list, bElem *arrA, bElem *arrB) {
  long size = 1;
  for (long n : list) size *= n;
  bool same = true;
#pragma omp parallel for reduction(&& : same)
  for (long l = 0; l < size; ++l) {
    const double diff = std::abs(arrA[l] - arrB[l]);
    same = same && ((diff < BRICK_TOLERANCE) || (diff < (std::abs(arrA[l]) + std::abs(arrB[l])) * BRICK_TOLERANCE));
  }
  return same;
}

#endif  // BRICKS_B200_MULTIARRAY_H
