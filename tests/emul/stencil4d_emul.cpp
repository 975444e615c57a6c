// TEST INFRASTRUCTURE.  Thread-by-thread host emulation of the 4-D kernels: the kernel bodies of
// bricklib_b200/csrc/stencil4d_body.cuh compiled as host code and run for every (block, thread) of the
// launch shape the library uses.  A thread of those kernels shares nothing with other threads, so this is
// the launch, minus the hardware: it checks slots, offsets, the marching registers and the launch shape
// where no GPU is present.  Not part of libbricks_b200.so and never a fallback.
#include "../../bricklib_b200/csrc/stencil4d_body.cuh"

extern "C" int emul_stencil4d(int which, const unsigned *grid, const unsigned *gdims, const unsigned *adj,
                              const double *in, double *out, unsigned long long in_step,
                              unsigned long long out_step, unsigned skip) {
  s4::Sweep4 p;
  p.grid = grid, p.adj = adj, p.in = in, p.out = out, p.in_step = in_step, p.out_step = out_step;
  p.n0 = gdims[0], p.n1 = gdims[1], p.n2 = gdims[2], p.n3 = gdims[3];
  p.s = skip, p.njg = p.nkg = 0;
  for (int a = 0; a < 4; ++a)
    if (gdims[a] <= 2 * skip) return 0;
  if (which == 0 || which >= 10) {
    /* which = 0: the default shape; else <LJ><LK> in decimal (22, 41, 21, 11) */
    const int lj = which ? which / 10 : 2, lk = which ? which % 10 : 2;
    const long long nb = (long long)s4::line_blocks(&p, lj, lk);
#pragma omp parallel for schedule(static)
    for (long long b = 0; b < nb; ++b)
      for (unsigned t = 0; t < (unsigned)(64 * lj * lk); ++t) {
        if (lj == 2 && lk == 2) s4::line_body<2, 2, false>(p, (unsigned)b, t);
        else if (lj == 4 && lk == 1) s4::line_body<4, 1, false>(p, (unsigned)b, t);
        else if (lj == 2 && lk == 1) s4::line_body<2, 1, false>(p, (unsigned)b, t);
        else s4::line_body<1, 1, false>(p, (unsigned)b, t);
      }
  } else {
    const long long nb = (long long)s4::gather_blocks(p);
#pragma omp parallel for schedule(static)
    for (long long b = 0; b < nb; ++b)
      for (unsigned t = 0; t < 256; ++t) s4::gather_body(p, (unsigned)b, t);
  }
  return 0;
}
