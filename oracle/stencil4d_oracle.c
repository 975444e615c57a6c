/*
 * oracle/stencil4d_oracle.c -- TEST INFRASTRUCTURE, not product code.
 *
 * Plain-C restatement of the reference's 4-D 9-point stencil on 4x4x4x4 bricks (the CPU-only path of
 * weak/main-4d.cpp).  It is the checker for bricks_stencil4d_* (bricklib_b200/csrc/stencil4d.cu); only
 * tests/, __graft_entry__.smoke() and bench.py's CPU legs may load it.
 * PINNED: tests/test_oracle_golden.py checks it against tests/golden/d4pt9_*.npz, produced by
 * oracle/gen_golden.py from the compiled, unmodified reference (oracle/_ref/libref4d.so: the codegen
 * kernel of stencils/mpi9pt.py with the vector settings weak/main-4d.cpp:36-58 selects).
 *
 * What it follows:
 *   - point set and coefficients: stencils/mpi9pt.py:12-27 -- alpha = 0.2 at the centre, then beta = 0.1
 *     at i+1, i-1, j+1, j-1, k+1, k-1, l+1, l-1 in that order (i = Index(0) is the contiguous axis);
 *   - neighbour access: include/brick.h:229-246 (_BrickAccessor::operator[]): per axis, index x in
 *     [-4, 8): dir = x + 4, neighbour digit dir / 4, in-brick index dir % 4; adjacency slot
 *     ((dl*3 + dk)*3 + dj)*3 + di with digit 1 = this brick; element
 *     dat[adj[b][slot]*step + ((l*4 + k)*4 + j)*4 + i] (row-major in-brick order, the device layout);
 *   - sweep region of one step: weak/main-4d.cpp:62-75 (brick_stencil: every grid cell from skip/TILE
 *     bricks inside the ghost-including grid on each side);
 *   - array form: weak/main-4d.cpp:171-193 (the eight neighbours summed l+1, l-1, k+1, k-1, j+1, j-1,
 *     i+1, i-1, times 0.1, plus the centre times 0.2).
 * Terms are summed one at a time, left to right, without fused multiply-add (-ffp-contract=off).
 */
#include <stddef.h>

#define B4 4
#define B4_ELEMS 256
#define B4_NADJ 81

static inline double at4(const unsigned *adj, const double *dat, long step, unsigned b, int l, int k, int j,
                         int i) {
  const int dl = l + B4, dk = k + B4, dj = j + B4, di = i + B4;
  const int slot = ((dl / B4 * 3 + dk / B4) * 3 + dj / B4) * 3 + di / B4;
  const unsigned nb = adj[(size_t)b * B4_NADJ + slot];
  return dat[(size_t)nb * step + (((dl % B4) * B4 + dk % B4) * B4 + dj % B4) * B4 + di % B4];
}

/* out[b] = stencil(in) for every brick b of blist; in and out are brick storages with `in_step` /
 * `out_step` elements per brick */
int oracle_brick_sweep_4d(const unsigned *adj, const unsigned *blist, long nlist, const double *in, double *out,
                          long in_step, long out_step) {
  const double alpha = 0.2, beta = 0.1;
#pragma omp parallel for schedule(static)
  for (long n = 0; n < nlist; ++n) {
    const unsigned b = blist[n];
    for (int l = 0; l < B4; ++l)
      for (int k = 0; k < B4; ++k)
        for (int j = 0; j < B4; ++j)
          for (int i = 0; i < B4; ++i) {
            double s = alpha * at4(adj, in, in_step, b, l, k, j, i);
            s += beta * at4(adj, in, in_step, b, l, k, j, i + 1);
            s += beta * at4(adj, in, in_step, b, l, k, j, i - 1);
            s += beta * at4(adj, in, in_step, b, l, k, j + 1, i);
            s += beta * at4(adj, in, in_step, b, l, k, j - 1, i);
            s += beta * at4(adj, in, in_step, b, l, k + 1, j, i);
            s += beta * at4(adj, in, in_step, b, l, k - 1, j, i);
            s += beta * at4(adj, in, in_step, b, l + 1, k, j, i);
            s += beta * at4(adj, in, in_step, b, l - 1, k, j, i);
            out[(size_t)b * out_step + ((l * B4 + k) * B4 + j) * B4 + i] = s;
          }
  }
  return 0;
}

/* brick list of one sweep of weak/main-4d.cpp:62-75: grid[tl][tk][tj][ti] for every axis index in
 * [s, n - s); gdims contiguous axis first; returns the count (blist may be NULL to size it) */
long oracle_sweep_list_4d(const unsigned *grid, const unsigned *gdims, unsigned s, unsigned *blist) {
  long n = 0;
  for (unsigned tl = s; tl + s < gdims[3]; ++tl)
    for (unsigned tk = s; tk + s < gdims[2]; ++tk)
      for (unsigned tj = s; tj + s < gdims[1]; ++tj)
        for (unsigned ti = s; ti + s < gdims[0]; ++ti) {
          if (blist) blist[n] = grid[((size_t)(tl * gdims[2] + tk) * gdims[1] + tj) * gdims[0] + ti];
          ++n;
        }
  return n;
}

/* array form over the box [lo, hi) per axis (contiguous axis first) of an array with extents ext[0..3] */
int oracle_array_sweep_4d(const double *in, double *out, const long *ext, const long *lo, const long *hi) {
  const long s1 = ext[0], s2 = ext[0] * ext[1], s3 = ext[0] * ext[1] * ext[2];
#pragma omp parallel for collapse(2) schedule(static)
  for (long l = lo[3]; l < hi[3]; ++l)
    for (long k = lo[2]; k < hi[2]; ++k)
      for (long j = lo[1]; j < hi[1]; ++j)
        for (long i = lo[0]; i < hi[0]; ++i) {
          const long p = l * s3 + k * s2 + j * s1 + i;
          double s = in[p + s3];
          s += in[p - s3];
          s += in[p + s2];
          s += in[p - s2];
          s += in[p + s1];
          s += in[p - s1];
          s += in[p + 1];
          s += in[p - 1];
          out[p] = s * 0.1 + in[p] * 0.2;
        }
  return 0;
}

/* copyToBrick<4> / copyFromBrick<4> (include/bricksetup.h:172-221) on row-major 4x4x4x4 bricks: the bricks
 * covering `dim` (+ `ghost` per side) of an array that carries pad + ghost extra elements per side */
void oracle_copy_brick_4d(int to_brick, const long *dim, const long *pad, const long *ghost, double *arr,
                          const unsigned *grid, double *bdat, long step) {
  long sA[4], sB[4], a = 1, b = 1, lo[4], hi[4];
  for (int d = 0; d < 4; ++d) {
    sA[d] = a;
    sB[d] = b;
    a *= dim[d] + 2 * (pad[d] + ghost[d]);
    b *= (dim[d] + 2 * ghost[d]) / B4;
    lo[d] = ghost[d] / B4;
    hi[d] = (dim[d] + ghost[d]) / B4;
  }
#pragma omp parallel for collapse(2)
  for (long tl = lo[3]; tl < hi[3]; ++tl)
    for (long tk = lo[2]; tk < hi[2]; ++tk)
      for (long tj = lo[1]; tj < hi[1]; ++tj)
        for (long ti = lo[0]; ti < hi[0]; ++ti) {
          const unsigned bi = grid[ti * sB[0] + tj * sB[1] + tk * sB[2] + tl * sB[3]];
          double *base = arr + (pad[0] + ti * B4) * sA[0] + (pad[1] + tj * B4) * sA[1] + (pad[2] + tk * B4) * sA[2] +
                         (pad[3] + tl * B4) * sA[3];
          for (int l = 0; l < B4; ++l)
            for (int k = 0; k < B4; ++k)
              for (int j = 0; j < B4; ++j)
                for (int i = 0; i < B4; ++i) {
                  double *e = bdat + (size_t)bi * step + ((l * B4 + k) * B4 + j) * B4 + i;
                  double *p = base + i * sA[0] + j * sA[1] + k * sA[2] + l * sA[3];
                  if (to_brick)
                    *e = *p;
                  else
                    *p = *e;
                }
        }
}
