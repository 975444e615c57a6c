/*
 * oracle/refshim/ref_stencil.cpp -- TEST INFRASTRUCTURE, not product code.
 *
 * extern "C" window onto the UNMODIFIED reference CPU stencil path: the
 * `brick(<stencil>.py, VSVEC, (8,8,8), (VFOLD), b)` call site of
 * weak/main.cpp:26-36 / stencils/3axis.cpp:68-76 expanded by the reference's own
 * codegen/vecscatter, and the literal array form ST_CPU of stencils/fake.h:41-352
 * (7-pt-with-coefficients: stencils/3axis.cpp:32-37).  oracle/Makefile runs the
 * reference code generator on this file and compiles it once per stencil
 * (-DREF_WHICH=7|13|25|125 for the MPI_* constant-coefficient stencils,
 * -DREF_WHICH=0 for stencils/7pt.py with the runtime `coeff[]` array, -DREF_WHICH=1 for
 * stencils/cond.py, the max/abs stencil of d3cond, stencils/3axis.cpp:94-163) into
 * oracle/_ref/.  Used by tests, oracle/gen_golden.py and bench.py's CPU baseline.
 *
 * fp32 bricks: built a second time with -DbElem=float -DREF_F32 (include/vecscatter.h:12-14 makes the
 * element type overridable; the code generator takes it from the pragma and emits 16-wide float
 * vectors).  Fold 2,8 = one AVX-512 float vector per two rows, which keeps a brick row-major.  The
 * symbols of that build are reff<N>_*; data pointers are `bElem *` throughout.
 */
#include <omp.h>
#include <algorithm>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <memory>

/* fake.h selects the stencil from MPI_25PT / MPI_13PT / MPI_125PT (else 7-pt). */
#if REF_WHICH == 13
#define MPI_13PT
#elif REF_WHICH == 25
#define MPI_25PT
#elif REF_WHICH == 125
#define MPI_125PT
#endif
#include "stencils/fake.h"
#include "stencils/cpuvfold.h"
#include <bricksetup.h>

#ifdef REF_F32
#ifndef __AVX512F__
#error "the fp32 reference build is AVX-512 only (16 floats per vector, fold 2,8)"
#endif
#undef VFOLD
#define VFOLD 2, 8
#endif
typedef bElem ref_elem;

#define REF_BDIM 8, 8, 8
typedef Brick<Dim<REF_BDIM>, Dim<VFOLD>> Brick3D;

#define REF_CAT2(a, b) a##b
#define REF_CAT(a, b) REF_CAT2(a, b)
#ifdef REF_F32
#define REFSYM(name) REF_CAT(REF_CAT(reff, REF_WHICH), _##name)
#else
#define REFSYM(name) REF_CAT(REF_CAT(ref, REF_WHICH), _##name)
#endif

#if REF_WHICH == 0 || REF_WHICH == 1
static bElem *coeff = nullptr; /* the free identifier stencils/7pt.py and cond.py name (one per translation unit) */
#endif
#if REF_WHICH == 1
using std::max; /* cond.py calls max(); stencils/3axis.cpp:12 */
#endif

namespace {

BrickStorage wrap(ref_elem *p, long chunks, long step) {
  BrickStorage s;
  s.dat = std::shared_ptr<bElem>(p, [](bElem *) {});
  s.chunks = chunks;
  s.step = (size_t)step;
  return s;
}

struct InfoView {
  BrickInfo<3> info;
  void *scratch;
  InfoView(const unsigned *adj, unsigned nbricks) : info(0) {
    scratch = info.adj;
    info.adj = (BrickInfo<3>::adjlist)adj;
    info.nbricks = nbricks;
  }
  ~InfoView() { free(scratch); }
};

/* offset of element (k,j,i) inside one brick in THIS build's fold order, obtained from the
 * reference accessor itself (include/brick.h:214-327). */
void fold_table(unsigned *tab) {
  unsigned adj1[27];
  for (int i = 0; i < 27; ++i) adj1[i] = 0;
  InfoView iv(adj1, 1);
  static ref_elem buf[512];
  BrickStorage st = wrap(buf, 1, 512);
  Brick3D b(&iv.info, st, 0);
  for (unsigned k = 0; k < 8; ++k)
    for (unsigned j = 0; j < 8; ++j)
      for (unsigned i = 0; i < 8; ++i) tab[k * 64 + j * 8 + i] = (unsigned)(&b[0][k][j][i] - buf);
}

}  // namespace

extern "C" {

int REFSYM(st_iter)() {
#if REF_WHICH == 0 || REF_WHICH == 1
  return 1;
#else
  return ST_ITER;
#endif
}

const char *REFSYM(vsvec)() { return VSVEC; }

void REFSYM(fold_table)(unsigned *tab512) { fold_table(tab512); }

/* One sweep of the codegen brick kernel over blist[0..nlist); storages are in the
 * reference's own fold order for this build (see fold_table). */
void REFSYM(brick_sweep)(const unsigned *adj, unsigned nbricks, const unsigned *blist, long nlist,
                         ref_elem *in_dat, ref_elem *out_dat, long step, const double *coeff_in) {
  InfoView iv(adj, nbricks);
  BrickStorage sin = wrap(in_dat, nbricks, step), sout = wrap(out_dat, nbricks, step);
#if REF_WHICH == 0 || REF_WHICH == 1
  coeff = const_cast<double *>(coeff_in);
  Brick3D bIn(&iv.info, sin, 0), bOut(&iv.info, sout, 0);
#else
  (void)coeff_in;
  Brick3D in(&iv.info, sin, 0), out(&iv.info, sout, 0);
#endif
#pragma omp parallel for schedule(static)
  for (long n = 0; n < nlist; ++n) {
    unsigned b = blist[n];
    brick(REF_PY, VSVEC, (REF_BDIM), (VFOLD), b);
  }
}

/* Same as brick_sweep but storages are row-major [k][j][i] bricks (the device layout);
 * converted through fold_table when this build folds differently. */
void REFSYM(brick_sweep_rm)(const unsigned *adj, unsigned nbricks, const unsigned *blist, long nlist,
                            const ref_elem *in_rm, ref_elem *out_rm, long step, const double *coeff_in) {
  unsigned tab[512];
  fold_table(tab);
  bool ident = true;
  for (unsigned e = 0; e < 512; ++e) ident = ident && (tab[e] == e);
  if (ident) {
    REFSYM(brick_sweep)(adj, nbricks, blist, nlist, const_cast<ref_elem *>(in_rm), out_rm, step, coeff_in);
    return;
  }
  size_t tot = (size_t)nbricks * step;
  ref_elem *fin = (ref_elem *)aligned_alloc(64, tot * sizeof(ref_elem));
  ref_elem *fout = (ref_elem *)aligned_alloc(64, tot * sizeof(ref_elem));
  long nfield = step / 512;
#pragma omp parallel for
  for (long c = 0; c < (long)nbricks * nfield; ++c)
    for (unsigned e = 0; e < 512; ++e) {
      fin[c * 512 + tab[e]] = in_rm[c * 512 + e];
      fout[c * 512 + tab[e]] = out_rm[c * 512 + e];
    }
  REFSYM(brick_sweep)(adj, nbricks, blist, nlist, fin, fout, step, coeff_in);
#pragma omp parallel for
  for (long c = 0; c < (long)nbricks * nfield; ++c)
    for (unsigned e = 0; e < 512; ++e) out_rm[c * 512 + e] = fout[c * 512 + tab[e]];
  free(fin);
  free(fout);
}

/* `reps` ping-pong sweeps (in->out, out->in, ...) timed with omp_get_wtime after one
 * warm-up sweep; returns seconds per sweep.  All host threads OpenMP gives us. */
double REFSYM(brick_sweep_timed)(const unsigned *adj, unsigned nbricks, const unsigned *blist,
                                 long nlist, ref_elem *a, ref_elem *b, long step, const double *coeff_in,
                                 int reps) {
  REFSYM(brick_sweep)(adj, nbricks, blist, nlist, a, b, step, coeff_in);
  double t0 = omp_get_wtime();
  for (int r = 0; r < reps; ++r) {
    if (r & 1)
      REFSYM(brick_sweep)(adj, nbricks, blist, nlist, b, a, step, coeff_in);
    else
      REFSYM(brick_sweep)(adj, nbricks, blist, nlist, a, b, step, coeff_in);
  }
  return (omp_get_wtime() - t0) / reps;
}

int REFSYM(max_threads)() { return omp_get_max_threads(); }

/* The reference's literal array stencil over the box [k0,k1) x [j0,j1) x [i0,i1) of a
 * padded array with extents (s0,s1,s2), contiguous dimension first. */
void REFSYM(array_sweep)(const ref_elem *in_ptr, ref_elem *out_ptr, long s0, long s1, long s2, long k0,
                         long k1, long j0, long j1, long i0, long i1, const double *coeff_in) {
  (void)s2;
  auto arrIn = (const bElem(*)[s1][s0])in_ptr;
  auto arrOut = (bElem(*)[s1][s0])out_ptr;
#if REF_WHICH == 0 || REF_WHICH == 1
  const double *c = coeff_in;
#else
  (void)coeff_in;
#endif
#pragma omp parallel for collapse(2)
  for (long k = k0; k < k1; ++k)
    for (long j = j0; j < j1; ++j)
#pragma omp simd
      for (long i = i0; i < i1; ++i)
#if REF_WHICH == 1
      {
        /* d3cond's array form, stencils/3axis.cpp:112-121: k+1, k-1, j+1, j-1, i+1, i-1, centre */
        const bElem partial = c[5] * max(arrIn[k + 1][j][i], 0.0) + c[6] * max(arrIn[k - 1][j][i], 0.0) +
                              c[3] * max(arrIn[k][j + 1][i], 0.0) + c[4] * max(arrIn[k][j - 1][i], 0.0) +
                              c[1] * max(arrIn[k][j][i + 1], 0.0) + c[2] * max(arrIn[k][j][i - 1], 0.0) +
                              c[0] * max(arrIn[k][j][i], 0.0);
        arrOut[k][j][i] = partial > 0 ? partial : -partial;
      }
#elif REF_WHICH == 0
        /* stencils/7pt.py:19-26 term order: centre, i+1, i-1, j+1, j-1, k+1, k-1 */
        arrOut[k][j][i] = c[0] * arrIn[k][j][i] + c[1] * arrIn[k][j][i + 1] +
                          c[2] * arrIn[k][j][i - 1] + c[3] * arrIn[k][j + 1][i] +
                          c[4] * arrIn[k][j - 1][i] + c[5] * arrIn[k + 1][j][i] +
                          c[6] * arrIn[k - 1][j][i];
#else
        ST_CPU;
#endif
}

}  // extern "C"
