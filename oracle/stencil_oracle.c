/*
 * oracle/stencil_oracle.c -- TEST INFRASTRUCTURE, not product code.
 *
 * Plain-C restatement of the reference stencil arithmetic on the brick layout.  It is the
 * checker for the CUDA kernels; only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline leg may load it.  PINNED: tests/test_oracle_golden.py checks it against
 * tests/golden/*.npz, which oracle/gen_golden.py produced from the compiled, unmodified
 * reference (oracle/_ref, i.e. the reference's codegen brick kernels and ST_CPU array forms).
 *
 * What it follows:
 *   - point sets + coefficients: stencils/7pt.py:19-26 (coeff[0..6]: centre, i+1, i-1, j+1,
 *     j-1, k+1, k-1), stencils/mpi7pt.py:17-24, mpi13pt.py:18-31, mpi25pt.py:20-45,
 *     mpi125pt.py:12-161 (10 constants by sorted |offset| class), constants stencils/fake.h:11-33
 *   - neighbour access semantics: include/brick.h:229-246 (_BrickAccessor::operator[]):
 *     for index x in [-8,16): dir = x + 8, neighbour digit dir / 8, in-brick index dir % 8,
 *     adjacency slot (dk*3 + dj)*3 + di, element dat[adj[b][slot]*step + k*64 + j*8 + i]
 *     (row-major in-brick order = fold <8> / <4,8>, include/brick.h:235-245)
 *   - array form: stencils/fake.h:41-352 (ST_CPU), stencils/3axis.cpp:32-37
 *   - d3pt27 (27 runtime coefficients): stencils/3axis.cpp:165-247; pinned against a block of the
 *     output of the reference's own d3pt27() run (oracle/refshim/ref_3axis.cpp,
 *     tests/golden/d3pt27_block.npz), term 16 read at (k+1, j+1, i) as the source has it
 * Terms are summed one at a time, left to right, in the DSL's order of coefficient classes,
 * without fused multiply-add (built with -ffp-contract=off).
 */
#include <stddef.h>
#include <stdlib.h>

#define MAXPTS 125

typedef struct {
  int n;
  int off[MAXPTS][3]; /* (di, dj, dk) */
  double c[MAXPTS];
} points_t;

static void add_pt(points_t *p, int di, int dj, int dk, double c) {
  p->off[p->n][0] = di;
  p->off[p->n][1] = dj;
  p->off[p->n][2] = dk;
  p->c[p->n] = c;
  p->n++;
}

/* star of radius R: centre then, per distance r = 1..R, i+r, i-r, j+r, j-r, k+r, k-r */
static void star(points_t *p, int R, const double *a) {
  add_pt(p, 0, 0, 0, a[0]);
  for (int r = 1; r <= R; ++r) {
    add_pt(p, r, 0, 0, a[r]);
    add_pt(p, -r, 0, 0, a[r]);
    add_pt(p, 0, r, 0, a[r]);
    add_pt(p, 0, -r, 0, a[r]);
    add_pt(p, 0, 0, r, a[r]);
    add_pt(p, 0, 0, -r, a[r]);
  }
}

static int cls125(int a, int b, int c) {
  /* sorted (|a|,|b|,|c|) -> class index of mpi125pt.py:12-33 */
  a = abs(a); b = abs(b); c = abs(c);
  int t;
  if (a > b) { t = a; a = b; b = t; }
  if (b > c) { t = b; b = c; c = t; }
  if (a > b) { t = a; a = b; b = t; }
  static const int tab[3][3][3] = {
      {{0, 1, 2}, {-1, 3, 4}, {-1, -1, 5}},
      {{-1, -1, -1}, {-1, 6, 7}, {-1, -1, 8}},
      {{-1, -1, -1}, {-1, -1, -1}, {-1, -1, 9}}};
  return tab[a][b][c];
}

/* which: 0 = 7pt.py (coeff_in[0..6]), 1 = cond.py (same, max/abs), 27 = d3pt27 (coeff_in[0..26]), 7/13/25/125 = mpi*pt.py constants */
int oracle_points(int which, const double *coeff_in, int *off_out, double *c_out) {
  points_t p;
  p.n = 0;
  if (which == 0 || which == 1) { /* 1 = cond.py: same point set, see oracle_is_cond */
    const double *c = coeff_in;
    add_pt(&p, 0, 0, 0, c[0]);
    add_pt(&p, 1, 0, 0, c[1]);
    add_pt(&p, -1, 0, 0, c[2]);
    add_pt(&p, 0, 1, 0, c[3]);
    add_pt(&p, 0, -1, 0, c[4]);
    add_pt(&p, 0, 0, 1, c[5]);
    add_pt(&p, 0, 0, -1, c[6]);
  } else if (which == 7) {
    const double a[2] = {0.4, 0.1}; /* MPI_ALPHA, MPI_BETA */
    star(&p, 1, a);
  } else if (which == 13) {
    const double a[3] = {0.4, 0.07, 0.03}; /* MPI_B0..B2 */
    star(&p, 2, a);
  } else if (which == 25) {
    const double a[5] = {0.1, 0.06, 0.045, 0.03, 0.015}; /* MPI_A0..A4 */
    star(&p, 4, a);
  } else if (which == 125) {
    const double C[10] = {0.1, 0.04, 0.03, 0.01, 0.006, 0.004, 0.005, 0.002, 0.003, 0.001};
    for (int cl = 0; cl < 10; ++cl)
      for (int dk = -2; dk <= 2; ++dk)
        for (int dj = -2; dj <= 2; ++dj)
          for (int di = -2; di <= 2; ++di)
            if (cls125(di, dj, dk) == cl) add_pt(&p, di, dj, dk, C[cl]);
  } else if (which == 27) {
    /* d3pt27 (stencils/3axis.cpp:186-200 array form, :209-231 brick form): term t multiplies
     * in[k + t/9 - 1][j + (t/3)%3 - 1][i + t%3 - 1] by coeff[t] -- except term 16, which BOTH forms
     * of the reference read at [k+1][j+1][i] instead of [k][j+1][i].  Restated as written. */
    const double *c = coeff_in;
    for (int t = 0; t < 27; ++t) {
      int dk = t / 9 - 1, dj = (t / 3) % 3 - 1, di = t % 3 - 1;
      if (t == 16) dk = 1;
      add_pt(&p, di, dj, dk, c[t]);
    }
  } else {
    return -1;
  }
  for (int n = 0; n < p.n; ++n) {
    if (off_out) {
      off_out[3 * n + 0] = p.off[n][0];
      off_out[3 * n + 1] = p.off[n][1];
      off_out[3 * n + 2] = p.off[n][2];
    }
    if (c_out) c_out[n] = p.c[n];
  }
  return p.n;
}

/* stencils/cond.py:23-30 (d3cond, stencils/3axis.cpp:94-163): every input goes through max(x, 0)
 * and the sum through If(calc > 0, calc, -calc) */
static inline int oracle_is_cond(int which) { return which == 1; }
static inline double cond_in(int cond, double v) { return cond && !(v > 0.0) ? 0.0 : v; }
static inline double cond_out(int cond, double v) { return cond && !(v > 0.0) ? -v : v; }

static inline double brick_at(const unsigned *adj, const double *dat, long step, unsigned b, int k,
                              int j, int i) {
  int dk = (k + 8) / 8, dj = (j + 8) / 8, di = (i + 8) / 8;
  int kk = (k + 8) % 8, jj = (j + 8) % 8, ii = (i + 8) % 8;
  unsigned nb = adj[(size_t)b * 27 + (dk * 3 + dj) * 3 + di];
  return dat[(size_t)nb * step + kk * 64 + jj * 8 + ii];
}

/* out[b] = stencil(in) for every b in blist; in/out are offsetted field pointers with their
 * own steps (interleaved fields: step 1024, offsets 0/512, stencils/3axis.cpp:25-28). */
int oracle_brick_sweep(int which, const unsigned *adj, const unsigned *blist, long nlist,
                       const double *in, long in_step, double *out, long out_step,
                       const double *coeff_in) {
  static int off[MAXPTS * 3];
  static double c[MAXPTS];
  int np = oracle_points(which, coeff_in, off, c);
  if (np < 0) return -1;
  const int cond = oracle_is_cond(which);
#pragma omp parallel for schedule(static)
  for (long n = 0; n < nlist; ++n) {
    unsigned b = blist[n];
    for (int k = 0; k < 8; ++k)
      for (int j = 0; j < 8; ++j)
        for (int i = 0; i < 8; ++i) {
          double acc = c[0] * cond_in(cond, brick_at(adj, in, in_step, b, k + off[2], j + off[1], i + off[0]));
          for (int t = 1; t < np; ++t)
            acc = acc + c[t] * cond_in(cond, brick_at(adj, in, in_step, b, k + off[3 * t + 2], j + off[3 * t + 1],
                                                      i + off[3 * t + 0]));
          out[(size_t)b * out_step + k * 64 + j * 8 + i] = cond_out(cond, acc);
        }
  }
  return 0;
}

/* fp32 bricks (bElem = float, include/vecscatter.h:12-14): the same point set and term order, the
 * coefficients rounded to float and every product and sum in float -- what the reference's generated
 * code computes when it is built with -DbElem=float (the constants of stencils/fake.h are double
 * literals, but the code generator emits them into float vectors).  Pinned by tests against that build
 * (oracle/_ref, reff<N>_brick_sweep) to a few ulp: summation order inside a vector differs. */
int oracle_brick_sweep_f32(int which, const unsigned *adj, const unsigned *blist, long nlist,
                           const float *in, long in_step, float *out, long out_step,
                           const double *coeff_in) {
  static int off[MAXPTS * 3];
  static double c[MAXPTS];
  int np = oracle_points(which, coeff_in, off, c);
  if (np < 0 || oracle_is_cond(which)) return -1;
#pragma omp parallel for schedule(static)
  for (long n = 0; n < nlist; ++n) {
    unsigned b = blist[n];
    for (int k = 0; k < 8; ++k)
      for (int j = 0; j < 8; ++j)
        for (int i = 0; i < 8; ++i) {
          float acc = 0.0f;
          for (int t = 0; t < np; ++t) {
            int kk = k + off[3 * t + 2], jj = j + off[3 * t + 1], ii = i + off[3 * t + 0];
            unsigned nb = adj[(size_t)b * 27 + (((kk + 8) / 8) * 3 + (jj + 8) / 8) * 3 + (ii + 8) / 8];
            float v = in[(size_t)nb * in_step + ((kk + 8) % 8) * 64 + ((jj + 8) % 8) * 8 + (ii + 8) % 8];
            float term = (float)c[t] * v;
            acc = t ? acc + term : term;
          }
          out[(size_t)b * out_step + k * 64 + j * 8 + i] = acc;
        }
  }
  return 0;
}

/* array twin on a padded array with extents (s0,s1,s2), contiguous first, over a box */
int oracle_array_sweep(int which, const double *in, double *out, long s0, long s1, long s2, long k0,
                       long k1, long j0, long j1, long i0, long i1, const double *coeff_in) {
  static int off[MAXPTS * 3];
  static double c[MAXPTS];
  (void)s2;
  int np = oracle_points(which, coeff_in, off, c);
  if (np < 0) return -1;
  const int cond = oracle_is_cond(which);
#pragma omp parallel for schedule(static)
  for (long k = k0; k < k1; ++k)
    for (long j = j0; j < j1; ++j)
      for (long i = i0; i < i1; ++i) {
        double acc = 0.0;
        for (int t = 0; t < np; ++t) {
          double v = cond_in(cond, in[((k + off[3 * t + 2]) * s1 + (j + off[3 * t + 1])) * s0 + i + off[3 * t]]);
          acc = t ? acc + c[t] * v : c[t] * v;
        }
        out[(k * s1 + j) * s0 + i] = cond_out(cond, acc);
      }
  return 0;
}

/* copyToBrick / copyFromBrick (include/bricksetup.h:139-221) for row-major bricks:
 * array extents per axis = dim + 2*(pad+ghost); brick grid extents = (dim+2*ghost)/8; the
 * loop covers grid cells [ghost/8, (dim+ghost)/8) per axis, array origin shifted by pad. */
void oracle_copy_brick(int to_brick, const long *dim, const long *pad, const long *ghost, double *arr,
                       const unsigned *grid, double *bdat, long step) {
  long sa[3], sb[3], na = 1, nb = 1;
  for (int a = 0; a < 3; ++a) {
    sa[a] = na;
    sb[a] = nb;
    na *= dim[a] + 2 * (pad[a] + ghost[a]);
    nb *= (dim[a] + 2 * ghost[a]) / 8;
  }
#pragma omp parallel for schedule(static)
  for (long bk = ghost[2] / 8; bk < (dim[2] + ghost[2]) / 8; ++bk)
    for (long bj = ghost[1] / 8; bj < (dim[1] + ghost[1]) / 8; ++bj)
      for (long bi = ghost[0] / 8; bi < (dim[0] + ghost[0]) / 8; ++bi) {
        unsigned b = grid[bk * sb[2] + bj * sb[1] + bi * sb[0]];
        for (int k = 0; k < 8; ++k)
          for (int j = 0; j < 8; ++j)
            for (int i = 0; i < 8; ++i) {
              double *a = arr + (pad[2] + bk * 8 + k) * sa[2] + (pad[1] + bj * 8 + j) * sa[1] +
                          (pad[0] + bi * 8 + i) * sa[0];
              double *e = bdat + (size_t)b * step + k * 64 + j * 8 + i;
              if (to_brick)
                *e = *a;
              else
                *a = *e;
            }
      }
}
