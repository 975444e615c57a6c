// Bodies of the 4-D kernels (stencil4d.cu), written so that the SAME text also compiles as host code: a
// thread of these kernels talks to no other thread (no shared memory, no barrier, no shuffle), so running
// body(block, thread) for every block and thread in a plain loop is an exact emulation of the launch.
// tests/emul/stencil4d_emul.cpp does that to check the index arithmetic and the marching logic against the
// oracle on machines without a GPU.  The emulation is test infrastructure: libbricks_b200.so contains only
// the __global__ wrappers.
#pragma once
#include <cstddef>

#ifdef __CUDACC__
#define S4_DEV __device__ __forceinline__
#else
#define S4_DEV inline
#endif

namespace s4 {

constexpr int kNAdj = 81;

struct Sweep4 {
  const unsigned *grid; /* [n3][n2][n1][n0] */
  const unsigned *adj;  /* nbricks x 81 */
  const double *in;
  double *out;
  unsigned long long in_step, out_step;
  unsigned n0, n1, n2, n3;
  unsigned s;        /* bricks skipped on each side */
  unsigned njg, nkg; /* CTA groups along j and k */
};

struct Row {
  double v0, v1, v2, v3;
};

S4_DEV unsigned ld_u32(const unsigned *p) {
#ifdef __CUDA_ARCH__
  return __ldg(p);
#else
  return *p;
#endif
}

S4_DEV double ld_f64(const double *p) {
#ifdef __CUDA_ARCH__
  return __ldg(p);
#else
  return *p;
#endif
}

/* one 32-byte row through the read-only path.  WIDE: one 256-bit load (LDG.E.256), so that a warp's 32 rows
 * are 8 full 128-byte lines per instruction; otherwise two 16-byte halves, each of which touches the same
 * 8 lines and uses half of every sector (twice the L1 wavefronts: profiles/r03_4d/ncu_full_line4d_128.txt). */
template <bool WIDE>
S4_DEV Row load_row(const double *base, unsigned long long step, unsigned brick, int off) {
  const double *q = base + (size_t)brick * step + off;
#ifdef __CUDA_ARCH__
  if (WIDE) {
    Row r;
    /* not volatile: the input is read-only for the whole launch, the scheduler may move the load freely */
    asm("ld.global.nc.v4.f64 {%0, %1, %2, %3}, [%4];" : "=d"(r.v0), "=d"(r.v1), "=d"(r.v2), "=d"(r.v3) : "l"(q));
    return r;
  }
  const double2 *p = reinterpret_cast<const double2 *>(q);
  const double2 a = __ldg(p), b = __ldg(p + 1);
  return Row{a.x, a.y, b.x, b.y};
#else
  return Row{q[0], q[1], q[2], q[3]};
#endif
}

/* one 32-byte streaming store */
S4_DEV void store_row(double *p, double a, double b, double c, double d) {
#ifdef __CUDA_ARCH__
  asm volatile("st.global.cs.v4.f64 [%0], {%1, %2, %3, %4};" ::"l"(p), "d"(a), "d"(b), "d"(c), "d"(d) : "memory");
#else
  p[0] = a, p[1] = b, p[2] = c, p[3] = d;
#endif
}

/* line_kernel: thread = one row position (l,k,j) of a line of bricks along the contiguous grid axis; a CTA
 * covers LJ x LK lines adjacent in j and k */
template <int LJ, int LK, bool WIDE>
S4_DEV void line_body(const Sweep4 &p, unsigned block, unsigned thread) {
  constexpr int kLJ = LJ, kLK = LK;
  const unsigned r = thread & 63; /* row (l,k,j) inside the brick */
  const unsigned line = thread >> 6;
  unsigned bid = block;
  const unsigned jg = bid % p.njg;
  bid /= p.njg;
  const unsigned kg = bid % p.nkg;
  const unsigned tl = p.s + bid / p.nkg;
  const unsigned tj = p.s + jg * kLJ + (line % kLJ);
  const unsigned tk = p.s + kg * kLK + (line / kLJ);
  if (tj >= p.n1 - p.s || tk >= p.n2 - p.s) return; /* no barrier below: whole warps may leave */

  const int j = r & 3, k = (r >> 2) & 3, l = r >> 4;
  const int roff = (int)r * 4;
  /* per axis a row sits on at most one face of its brick: that face's adjacency slot (40 = this brick) */
  const unsigned sj = j == 0 ? 37u : (j == 3 ? 43u : 40u);
  const unsigned sk = k == 0 ? 31u : (k == 3 ? 49u : 40u);
  const unsigned sl = l == 0 ? 13u : (l == 3 ? 67u : 40u);
  /* element offsets of the six neighbour rows: inside this brick, or wrapped into the neighbour */
  const int o_jm = roff + (j == 0 ? 12 : -4), o_jp = roff + (j == 3 ? -12 : 4);
  const int o_km = roff + (k == 0 ? 48 : -16), o_kp = roff + (k == 3 ? -48 : 16);
  const int o_lm = roff + (l == 0 ? 192 : -64), o_lp = roff + (l == 3 ? -192 : 64);

  const unsigned *gline = p.grid + ((size_t)(tl * p.n2 + tk) * p.n1 + tj) * p.n0;
  const unsigned lo = p.s, hi = p.n0 - p.s;
  const double alpha = 0.2, beta = 0.1; /* stencils/mpi9pt.py:12-13 */

  unsigned b = ld_u32(gline + lo);
  const unsigned *arow = p.adj + (size_t)b * kNAdj;
  unsigned aj = ld_u32(arow + sj), ak = ld_u32(arow + sk), al = ld_u32(arow + sl);
  unsigned a39 = ld_u32(arow + 39), a41 = ld_u32(arow + 41);
  Row c = load_row<WIDE>(p.in, p.in_step, b, roff);
  double prev3 = ld_f64(p.in + (size_t)a39 * p.in_step + roff + 3);

  for (unsigned ti = lo;; ++ti) {
    /* ids of the next step first: grid -> adj -> rows is the longest dependency chain, so it is in flight
     * while this step loads and computes.  The last step re-reads its own ids (no branch to sink them under). */
    const bool more = ti + 1 < hi;
    const unsigned bn = ld_u32(gline + (more ? ti + 1 : ti));
    const unsigned *nrow = p.adj + (size_t)bn * kNAdj;
    const unsigned naj = ld_u32(nrow + sj), nak = ld_u32(nrow + sk), nal = ld_u32(nrow + sl);
    const unsigned na39 = ld_u32(nrow + 39), na41 = ld_u32(nrow + 41);
    /* the i+ neighbour's row: its first element is this step's halo, the whole row the next centre */
    const Row nx = load_row<WIDE>(p.in, p.in_step, a41, roff);
    const Row jm = load_row<WIDE>(p.in, p.in_step, j == 0 ? aj : b, o_jm);
    const Row jp = load_row<WIDE>(p.in, p.in_step, j == 3 ? aj : b, o_jp);
    const Row km = load_row<WIDE>(p.in, p.in_step, k == 0 ? ak : b, o_km);
    const Row kp = load_row<WIDE>(p.in, p.in_step, k == 3 ? ak : b, o_kp);
    const Row lm = load_row<WIDE>(p.in, p.in_step, l == 0 ? al : b, o_lm);
    const Row lp = load_row<WIDE>(p.in, p.in_step, l == 3 ? al : b, o_lp);
    /* term order of stencils/mpi9pt.py:17-26: centre, i+1, i-1, j+1, j-1, k+1, k-1, l+1, l-1 */
    double o0 = alpha * c.v0, o1 = alpha * c.v1, o2 = alpha * c.v2, o3 = alpha * c.v3;
    o0 += beta * c.v1, o1 += beta * c.v2, o2 += beta * c.v3, o3 += beta * nx.v0;
    o0 += beta * prev3, o1 += beta * c.v0, o2 += beta * c.v1, o3 += beta * c.v2;
    o0 += beta * jp.v0, o1 += beta * jp.v1, o2 += beta * jp.v2, o3 += beta * jp.v3;
    o0 += beta * jm.v0, o1 += beta * jm.v1, o2 += beta * jm.v2, o3 += beta * jm.v3;
    o0 += beta * kp.v0, o1 += beta * kp.v1, o2 += beta * kp.v2, o3 += beta * kp.v3;
    o0 += beta * km.v0, o1 += beta * km.v1, o2 += beta * km.v2, o3 += beta * km.v3;
    o0 += beta * lp.v0, o1 += beta * lp.v1, o2 += beta * lp.v2, o3 += beta * lp.v3;
    o0 += beta * lm.v0, o1 += beta * lm.v1, o2 += beta * lm.v2, o3 += beta * lm.v3;
    store_row(p.out + (size_t)b * p.out_step + roff, o0, o1, o2, o3);
    if (!more) break;
    /* the grid's next brick is normally this brick's i+ neighbour and this brick its i- neighbour: the rows
     * are already in registers.  Anything else (a grid that is not a lattice of adj) reloads them.  Both
     * tests are uniform over the 64 threads of a brick. */
    const double c3 = c.v3;
    c = (bn == a41) ? nx : load_row<WIDE>(p.in, p.in_step, bn, roff);
    prev3 = (na39 == b) ? c3 : ld_f64(p.in + (size_t)na39 * p.in_step + roff + 3);
    b = bn, aj = naj, ak = nak, al = nal, a39 = na39, a41 = na41;
  }
}

/* gather_kernel (verification twin): one CTA per brick of the sweep box, one thread per point, every access
 * through the accessor arithmetic of include/brick.h:229-246 */
S4_DEV void gather_body(const Sweep4 &p, unsigned block, unsigned thread) {
  unsigned bid = block;
  const unsigned w0 = p.n0 - 2 * p.s, w1 = p.n1 - 2 * p.s, w2 = p.n2 - 2 * p.s;
  const unsigned ti = p.s + bid % w0;
  bid /= w0;
  const unsigned tj = p.s + bid % w1;
  bid /= w1;
  const unsigned tk = p.s + bid % w2;
  const unsigned tl = p.s + bid / w2;
  const unsigned b = p.grid[((size_t)(tl * p.n2 + tk) * p.n1 + tj) * p.n0 + ti];
  const unsigned *row = p.adj + (size_t)b * kNAdj;
  const int i = thread & 3, j = (thread >> 2) & 3, k = (thread >> 4) & 3, l = thread >> 6;
  auto at = [&](int dl, int dk, int dj, int di) -> double {
    const int ll = l + dl + 4, kk = k + dk + 4, jj = j + dj + 4, ii = i + di + 4;
    const unsigned nb = row[(((ll >> 2) * 3 + (kk >> 2)) * 3 + (jj >> 2)) * 3 + (ii >> 2)];
    return p.in[(size_t)nb * p.in_step + (((ll & 3) * 4 + (kk & 3)) * 4 + (jj & 3)) * 4 + (ii & 3)];
  };
  const double alpha = 0.2, beta = 0.1;
  double acc = alpha * at(0, 0, 0, 0);
  acc += beta * at(0, 0, 0, 1);
  acc += beta * at(0, 0, 0, -1);
  acc += beta * at(0, 0, 1, 0);
  acc += beta * at(0, 0, -1, 0);
  acc += beta * at(0, 1, 0, 0);
  acc += beta * at(0, -1, 0, 0);
  acc += beta * at(1, 0, 0, 0);
  acc += beta * at(-1, 0, 0, 0);
  p.out[(size_t)row[40] * p.out_step + thread] = acc;
}

/* launch shapes, shared by the launcher and the emulation */
inline unsigned long long line_blocks(Sweep4 *p, int lj, int lk) {
  const unsigned w1 = p->n1 - 2 * p->s, w2 = p->n2 - 2 * p->s, w3 = p->n3 - 2 * p->s;
  p->njg = (w1 + lj - 1) / lj;
  p->nkg = (w2 + lk - 1) / lk;
  return (unsigned long long)p->njg * p->nkg * w3;
}
inline unsigned long long gather_blocks(const Sweep4 &p) {
  return (unsigned long long)(p.n0 - 2 * p.s) * (p.n1 - 2 * p.s) * (p.n2 - 2 * p.s) * (p.n3 - 2 * p.s);
}

}  // namespace s4
