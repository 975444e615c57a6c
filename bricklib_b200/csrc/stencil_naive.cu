// Verification twin of the marching kernels: one thread per point straight through the
// adjacency list, the shape of the reference's d3pt7_brick (stencils/3axis.cu:12-26) and of the
// accessor semantics in include/brick.h:229-246.  Used by tests to triangulate a disagreement
// between the marching kernel and the CPU oracle; the product path (bricks_stencil) never
// calls it.
#include "common.cuh"

namespace {

template <typename T>
struct NaiveParams {
  const unsigned *adj;
  const unsigned *blist;
  const T *in;
  T *out;
  unsigned long long in_step, out_step;
  int npts;
  int cond; /* cond.py: max(in, 0) per term, |sum| */
};

__constant__ int c_off[125 * 3];
__constant__ double c_coef[125];

/* T = double, or float for fp32 bricks (bElem = float, include/vecscatter.h:12-14): same point set, the
 * coefficients rounded to T, every term accumulated in T as the reference's generated code does */
template <typename T>
__global__ void __launch_bounds__(512) naive_kernel(NaiveParams<T> p) {
  const unsigned b = p.blist[blockIdx.x];
  const int i = threadIdx.x & 7, j = (threadIdx.x >> 3) & 7, k = threadIdx.x >> 6;
  const unsigned *row = p.adj + (size_t)b * 27;
  T acc = 0;
  for (int t = 0; t < p.npts; ++t) {
    const int ii = i + c_off[3 * t] + 8, jj = j + c_off[3 * t + 1] + 8, kk = k + c_off[3 * t + 2] + 8;
    const unsigned nb = row[((kk >> 3) * 3 + (jj >> 3)) * 3 + (ii >> 3)];
    T v = p.in[(size_t)nb * p.in_step + (kk & 7) * 64 + (jj & 7) * 8 + (ii & 7)];
    if (p.cond) v = fmax(v, (T)0);
    const T c = (T)c_coef[t];
    acc = t ? acc + c * v : c * v;
  }
  p.out[(size_t)b * p.out_step + threadIdx.x] = p.cond ? (T)fabs(acc) : acc;
}

int build_points(int id, const double *coeff, int *off, double *cf) {
  int n = 0;
  auto add = [&](int di, int dj, int dk, double c) {
    off[3 * n] = di, off[3 * n + 1] = dj, off[3 * n + 2] = dk, cf[n] = c;
    ++n;
  };
  auto star = [&](int R, const double *a) {
    add(0, 0, 0, a[0]);
    for (int r = 1; r <= R; ++r) {
      add(r, 0, 0, a[r]), add(-r, 0, 0, a[r]);
      add(0, r, 0, a[r]), add(0, -r, 0, a[r]);
      add(0, 0, r, a[r]), add(0, 0, -r, a[r]);
    }
  };
  if (id == BRICKS_ST_7PT_COEFF || id == BRICKS_ST_COND) {
    if (!coeff) return -1;
    add(0, 0, 0, coeff[0]), add(1, 0, 0, coeff[1]), add(-1, 0, 0, coeff[2]);
    add(0, 1, 0, coeff[3]), add(0, -1, 0, coeff[4]), add(0, 0, 1, coeff[5]), add(0, 0, -1, coeff[6]);
  } else if (id == BRICKS_ST_27PT_COEFF) {
    if (!coeff) return -1;
    /* stencils/3axis.cpp:186-231, including term 16 read at [k+1][j+1][i] as the reference has it */
    for (int t = 0; t < 27; ++t) add(t % 3 - 1, (t / 3) % 3 - 1, t == 16 ? 1 : t / 9 - 1, coeff[t]);
  } else if (id == BRICKS_ST_7PT) {
    const double a[] = {0.4, 0.1};
    star(1, a);
  } else if (id == BRICKS_ST_13PT) {
    const double a[] = {0.4, 0.07, 0.03};
    star(2, a);
  } else if (id == BRICKS_ST_25PT) {
    const double a[] = {0.1, 0.06, 0.045, 0.03, 0.015};
    star(4, a);
  } else if (id == BRICKS_ST_125PT) {
    const double C[10] = {0.1, 0.04, 0.03, 0.01, 0.006, 0.004, 0.005, 0.002, 0.003, 0.001};
    const int cls[3][3][3] = {{{0, 1, 2}, {1, 3, 4}, {2, 4, 5}},
                              {{1, 3, 4}, {3, 6, 7}, {4, 7, 8}},
                              {{2, 4, 5}, {4, 7, 8}, {5, 8, 9}}};
    for (int dk = -2; dk <= 2; ++dk)
      for (int dj = -2; dj <= 2; ++dj)
        for (int di = -2; di <= 2; ++di) add(di, dj, dk, C[cls[abs(dk)][abs(dj)][abs(di)]]);
  } else {
    return -1;
  }
  return n;
}

template <typename T>
int run_naive(bricks_plan *plan, int stencil_id, int nfields, const T *const *in, T *const *out, size_t in_step,
              size_t out_step, const double *coeff_host, int stream) {
  BRK_ARG(plan && in && out && nfields > 0 && stream >= 0 && stream < bricks::kNumStreams);
  BRK_CUDA(cudaSetDevice(plan->ctx->device));
  cudaStream_t s = plan->ctx->streams[stream];
  int off[125 * 3];
  double cf[125];
  const int n = build_points(stencil_id, coeff_host, off, cf);
  if (n < 0) {
    bricks::set_error("unknown stencil id %d (or missing coefficients)", stencil_id);
    return -1;
  }
  BRK_CUDA(cudaMemcpyToSymbolAsync(c_off, off, sizeof(int) * 3 * n, 0, cudaMemcpyHostToDevice, s));
  BRK_CUDA(cudaMemcpyToSymbolAsync(c_coef, cf, sizeof(double) * n, 0, cudaMemcpyHostToDevice, s));
  BRK_CUDA(cudaStreamSynchronize(s)); /* off/cf are stack arrays */
  for (int f = 0; f < nfields; ++f) {
    NaiveParams<T> p;
    p.adj = plan->adj_dev;
    p.blist = plan->blist_dev;
    p.in = in[f];
    p.out = out[f];
    p.in_step = in_step;
    p.out_step = out_step;
    p.npts = n;
    p.cond = stencil_id == BRICKS_ST_COND;
    naive_kernel<T><<<(unsigned)plan->nlist, 512, 0, s>>>(p);
    BRK_CUDA(cudaGetLastError());
    ++bricks::g_launches;
  }
  return 0;
}

}  // namespace

extern "C" int bricks_stencil_naive(bricks_plan *plan, int stencil_id, int nfields,
                                    const double *const *in, double *const *out, size_t in_step,
                                    size_t out_step, const double *coeff_host, int stream) {
  return run_naive<double>(plan, stencil_id, nfields, in, out, in_step, out_step, coeff_host, stream);
}

extern "C" int bricks_stencil_naive_f32(bricks_plan *plan, int stencil_id, int nfields,
                                        const float *const *in, float *const *out, size_t in_step,
                                        size_t out_step, const double *coeff_host, int stream) {
  return run_naive<float>(plan, stencil_id, nfields, in, out, in_step, out_step, coeff_host, stream);
}
