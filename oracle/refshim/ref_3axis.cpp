/*
 * oracle/refshim/ref_3axis.cpp -- TEST INFRASTRUCTURE, not product code.
 *
 * Runs the reference's own single-node driver function d3pt27() (stencils/3axis.cpp:165-247:
 * the 27-point stencil with 27 runtime coefficients, array form and brick form, N = 256) from the
 * UNMODIFIED source file, and keeps the arrays it would free at its end so that
 * oracle/gen_golden.py can cut a small golden block out of them.
 *
 * d3pt27() owns its arrays (random input from std::random_device, so nothing about them is
 * reproducible from outside) and frees them before it returns.  The only hook is the preprocessor:
 * `free(p)` is redirected to ref_keep(p) for the text of 3axis.cpp and of the reference headers it
 * pulls in (the BrickStorage deleter of include/brick.h:72 included), after every standard header
 * those files need has already been included here.  Order of the kept pointers at return, from the
 * source: in_ptr, out_ptr, grid_ptr, bInfo.adj, then the brick storage (released when the
 * function's locals go out of scope).  d3pt27() itself throws when its brick result and its array
 * result differ (compareBrick), so a normal return also says the reference's two forms agree.
 *
 * Built by `make -C oracle golden27` only (it is not part of libref_*.so: one run takes ~30 s and
 * ~1 GB); nothing here is used at test time.
 */
#include <omp.h>
#include <stdlib.h>
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <initializer_list>
#include <iostream>
#include <memory>
#include <random>
#include <stdexcept>
#include <type_traits>
#include <typeinfo>
#include <vector>
#include <immintrin.h>

extern "C" void ref_keep(void *p);
#define free(p) ref_keep(p)
#include "stencils/3axis.cpp"
#undef free

bElem *coeff = nullptr; /* stencils/stencils.h:75; the reference defines it in stencils/main.cpp */

namespace {
std::vector<void *> kept;
}

extern "C" void ref_keep(void *p) { kept.push_back(p); }

extern "C" {

/* geometry of the arrays d3pt27() works on (stencils/stencils.h:11-25) */
void ref3axis_geometry(long *n, long *stride, long *first) {
  *n = N;
  *stride = STRIDE;
  *first = GZ + PADDING; /* index of the first computed point on every axis */
}

/* run d3pt27() with these 27 coefficients; returns 0 and the input / output arrays
 * (STRIDE^3 doubles each, [k][j][i]) on success, 1 when the reference reported a mismatch between
 * its brick and array results, 2 when the bookkeeping of kept pointers is not what the source
 * promises */
int ref3axis_d3pt27(const double *c27, double **in, double **out) {
  static double c[27];
  for (int i = 0; i < 27; ++i) c[i] = c27[i];
  coeff = c;
  kept.clear();
  try {
    d3pt27();
  } catch (const std::exception &e) {
    fprintf(stderr, "ref3axis_d3pt27: %s\n", e.what());
    return 1;
  }
  if (kept.size() != 5) return 2;
  *in = (double *)kept[0];
  *out = (double *)kept[1];
  return 0;
}

void ref3axis_release(void) {
  for (void *p : kept) free(p);
  kept.clear();
}

}  // extern "C"
