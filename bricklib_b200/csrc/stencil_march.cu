// The marching stencil kernels and their launcher (bricks_stencil).
// Replaces the generated bodies of `brick(ST_SCRTPT, "CUDA", (8,8,8), (4,8), b)` inside
// brick_kernel (weak/main.cu:35-43, strong/main.cu:85-111) and d3pt7_brick_trans
// (stencils/3axis.cu:28-37).  See march_dev.cuh for the work decomposition.
#include <cstdlib>
#include <cstring>

#include "march_steps.cuh"

/* Register caps.  The register file is 4 x 16 Ki, one quarter per scheduler, so a one-warp CTA fits
 * floor(16384 / (32 * regs)) times per scheduler: <= 128 registers -> 16 warps per SM, <= 168 ->
 * 12, more -> 8 (ncu launch__occupancy_limit_registers).  7-point: 128; 13-point and cube: 168. */
#ifndef MARCH_STAR4_REGS /* 25-point star: 216-230 registers as compiled, 8 warps per SM */
#define MARCH_STAR4_REGS 255
#endif
#ifndef MARCH_CUBE_REGS
#define MARCH_CUBE_REGS 168
#endif

#ifndef MARCH_F32_CUBE_REGS /* float kernels: 145 / 158 registers as compiled (12 warps per SM); capped at 128 they have no spills and 16 warps: cube 0.441 -> 0.427 ms, 25-point 0.346 -> 0.342 ms, profiles/r03_fp32_kernels.log */
#define MARCH_F32_CUBE_REGS 128
#endif
#ifndef MARCH_F32_STAR4_REGS
#define MARCH_F32_STAR4_REGS 128
#endif

namespace dev {
template <int R, bool CUBE, size_t EB>
constexpr int reg_cap() {
  if (EB == 8) return CUBE ? MARCH_CUBE_REGS : (R == 1 ? 128 : (R == 2 ? 168 : MARCH_STAR4_REGS));
  return CUBE ? (R == 2 ? MARCH_F32_CUBE_REGS : 128) : (R == 4 ? MARCH_F32_STAR4_REGS : 128);
}
}  // namespace dev

namespace {

template <int R, bool CUBE, int VAR, int P, typename T>
/* register cap for the cube: measured best on B200 (profiles/r01_maxnreg_sweep.txt) */
__global__ void __launch_bounds__(32) __maxnreg__((dev::reg_cap<R, CUBE, sizeof(T)>())) march_kernel(const __grid_constant__ MarchParams prm) {
  using L = dev::StageLayout<R, CUBE, P, T>;
  constexpr int H = 8 / P;
  extern __shared__ __align__(1024) char smem[];
  const int lane = threadIdx.x;
  const MarchItem it = prm.items[blockIdx.x];
  const FieldDev &fd = prm.fields[blockIdx.y];
  const int nk = it.nk, nsteps = nk * H + 2;

  const uint32_t smem0 = dev::smem_u32(smem);
  if ((smem0 & 511u) != 0) __trap(); /* swizzled centre pieces need 512-byte alignment */
  const uint32_t bar0 = smem0 + 2 * L::BYTES;
  if (lane == 0) {
    dev::mbar_init(bar0, 1);
    dev::mbar_init(bar0 + 8, 1);
    dev::fence_barrier_init();
    dev::fence_proxy_async();
  }
  __syncwarp();

  /* brick ids of the column: lane t holds layer t */
  unsigned col0 = 0, col1 = 0;
  if (lane < nk) {
    const size_t cell = ((size_t)(it.gk0 + lane) * prm.gd1 + it.gj) * prm.gd0 + it.gi;
    col0 = prm.grid[cell];
    col1 = it.valid1 ? prm.grid[cell + 1] : prm.adj[(size_t)col0 * 27 + 14];
  }

  dev::LaneK<R, CUBE, P, T> lk;
  lk.init(lane);
  const dev::PieceK pk = dev::piece_of<R, CUBE, P, T>(lane < 12 ? lane : 0);
  constexpr int MAIN_PLANE_BYTES =
      2 * L::C_PL + 2 * L::I_PL + L::JSUM_PL + (CUBE ? 4 * L::X_PL : 0);
  constexpr int ROLL_PLANE_BYTES = CUBE ? MAIN_PLANE_BYTES : 2 * L::C_PL;

  auto step_layer = [&](int s, int &dk) -> int {
    if (s == 0) { dk = -1; return 0; }
    if (s == nsteps - 1) { dk = +1; return nk - 1; }
    dk = 0;
    return (s - 1) / H;
  };
  /* neighbour brick that lane's piece of stage s is cut from */
  auto load_nb = [&](int s) -> unsigned {
    int dk;
    const int layer = step_layer(s, dk);
    const unsigned b0 = __shfl_sync(0xffffffffu, col0, layer);
    const unsigned b1 = __shfl_sync(0xffffffffu, col1, layer);
    const unsigned rowb = pk.useb1 ? b1 : b0;
    unsigned nb = 0;
    if (lane < 12) nb = prm.adj[(size_t)rowb * 27 + (dk + 1) * 9 + (pk.dj + 1) * 3 + (pk.di + 1)];
    return nb;
  };
  auto issue = [&](int s, unsigned nb) {
    const bool roll = (s == 0) || (s == nsteps - 1);
    const uint32_t bar = bar0 + 8 * (s & 1);
    const uint32_t dst = smem0 + (s & 1) * L::BYTES + pk.off;
    const int ck = (s == 0) ? 8 - R : (roll ? 0 : ((s - 1) % H) * P);
    if (lane == 0) dev::mbar_expect_tx(bar, roll ? R * ROLL_PLANE_BYTES : P * MAIN_PLANE_BYTES);
    const int np = roll ? L::NPIECE_ROLL : L::NPIECE_MAIN;
    /* pieces 0/1 (centres), 4/5 and 6/7 (j-faces) sit back to back in the stage: when their two
     * bricks are consecutive in storage -- the rule inside a BrickDecomp region -- the even lane
     * fills both with one two-brick copy and the odd lane stays idle (9 instead of 12 copies per
     * cube stage: +1-2 %; going further down to 5 changed nothing, profiles/r02_experiments/). */
    const unsigned nb_up = __shfl_down_sync(0xffffffffu, nb, 1), nb_dn = __shfl_up_sync(0xffffffffu, nb, 1);
    const bool pairable = !roll && (lane < 2 || (lane >= 4 && lane < 8));
    const bool lead = pairable && !(lane & 1) && nb_up == nb + 1;
    const bool follow = pairable && (lane & 1) && nb == nb_dn + 1;
    if (lane < np && !follow) {
      const CUtensorMap *map = roll ? &fd.maps.roll[pk.map] : (lead ? &fd.maps.pair[pk.pairmap] : &fd.maps.main[pk.map]);
      dev::tma_load_4d(dst, map, pk.ci, pk.cj, ck, (int)nb, bar);
    }
  };

  T carry[2 * R][4];
#pragma unroll
  for (int a = 0; a < 2 * R; ++a)
#pragma unroll
    for (int m = 0; m < 4; ++m) carry[a][m] = 0.0;

  dev::OutPtrs<T> o;
  o.cur = o.prev = nullptr;
  o.curv = o.prevv = false;
  const int lane_off = lk.j * 8 + lk.ih * 4;
  /* read once: behind the asm memory clobbers of the pipeline the compiler would re-read
   * fd.out / fd.out_step from global memory every layer (LDG.128 + a long-scoreboard wait in front
   * of the first store: 125-point 0.575 -> 0.562 ms, 25-point 0.629 -> 0.613 ms, the other kernels
   * unchanged; profiles/r02_experiments/variants_h_*, variants_i_*) */
  T *const out_base = (T *)fd.out;
  const unsigned long long out_stride = fd.out_step;

  /* coefficients in the element type: fp32 bricks compute in float as the reference's generated code does */
  const T *cc;
  T cf[27];
  if constexpr (sizeof(T) == 8) {
    cc = prm.c;
  } else {
#pragma unroll
    for (int i = 0; i < 27; ++i) cf[i] = (T)prm.c[i];
    cc = cf;
  }

  unsigned nbn = load_nb(0);
  issue(0, nbn);
  nbn = load_nb(1);

  for (int s = 0; s < nsteps; ++s) {
    if (s + 1 < nsteps) {
      dev::fence_proxy_async();
      issue(s + 1, nbn);
      if (s + 2 < nsteps) nbn = load_nb(s + 2);
    }
    dev::mbar_wait(bar0 + 8 * (s & 1), (s >> 1) & 1);
    const char *st = smem + (s & 1) * L::BYTES;
    int mode = 1, zbase = 0;
    if (s == 0) {
      mode = 0;
    } else if (s == nsteps - 1) {
      mode = 2;
      zbase = 8;
    } else {
      const int layer = (s - 1) / H, half = (s - 1) % H;
      zbase = half * P;
      if (half == 0) {
        o.prev = o.cur;
        o.prevv = o.curv && layer > 0;
        const unsigned b0 = __shfl_sync(0xffffffffu, col0, layer);
        const unsigned b1 = __shfl_sync(0xffffffffu, col1, layer);
        o.cur = out_base + (size_t)(lk.bb ? b1 : b0) * out_stride + lane_off;
        o.curv = (lk.bb == 0) || (it.valid1 != 0);
      }
    }
    if constexpr (CUBE && R == 1) {
      if (mode == 1)
        dev::cube27_step<P, P, 1, T>(st, lk, cc, carry, o, zbase);
      else if (mode == 0)
        dev::cube27_step<P, R, 0, T>(st, lk, cc, carry, o, zbase);
      else
        dev::cube27_step<P, R, 2, T>(st, lk, cc, carry, o, zbase);
    } else if constexpr (CUBE) {
      if (mode == 1)
        dev::cube_step<P, P, 1, T>(st, lk, cc, carry, o, zbase);
      else if (mode == 0)
        dev::cube_step<P, R, 0, T>(st, lk, cc, carry, o, zbase);
      else
        dev::cube_step<P, R, 2, T>(st, lk, cc, carry, o, zbase);
    } else {
      if (mode == 1)
        dev::star_step<R, VAR, P, P, 1, T>(st, lk, cc, carry, o, zbase);
      else if (mode == 0)
        dev::star_step<R, VAR, P, R, 0, T>(st, lk, cc, carry, o, zbase);
      else
        dev::star_step<R, VAR, P, R, 2, T>(st, lk, cc, carry, o, zbase);
    }
    __syncwarp();
  }
}

template <int R, bool CUBE, int VAR, int P, typename T>
int launch(bricks_plan *plan, int stencil_id, const FieldDev *fields, int nfields, const double *c, cudaStream_t s) {
  using L = dev::StageLayout<R, CUBE, P, T>;
  const int smem = 2 * L::BYTES + 64;
  auto kern = march_kernel<R, CUBE, VAR, P, T>;
  /* per device, once; host threads that drive different devices may be here concurrently (setting
   * the attributes twice is harmless, the mask only saves the calls) */
  static std::atomic<unsigned long long> configured{0};
  const unsigned long long bit = 1ull << (plan->ctx->device & 63);
  if (!(configured.load(std::memory_order_acquire) & bit)) {
    BRK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    BRK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
    configured.fetch_or(bit, std::memory_order_release);
  }
  MarchParams prm;
  int nitems = 0;
  if (int rc = bricks::plan_items(plan, bricks::column_length(stencil_id), &prm.items, &nitems)) return rc;
  prm.grid = plan->grid_dev;
  prm.adj = plan->adj_dev;
  prm.fields = fields;
  prm.gd0 = plan->gdims[0];
  prm.gd1 = plan->gdims[1];
  for (int i = 0; i < 27; ++i) prm.c[i] = c[i];
  kern<<<dim3(nitems, nfields), 32, smem, s>>>(prm);
  BRK_CUDA(cudaGetLastError());
  ++bricks::g_launches;
  return 0;
}

/* planes of a brick layer per pipeline stage: two stages of four planes (measured against eight,
 * profiles/r01_tma_piece_experiments.txt: fewer one-warp CTAs fit) */
constexpr int kPlanes = 4;

/* build (or find) the device FieldDev array for these storages and this kernel variant */
template <typename T>
int get_fields(bricks_plan *plan, int stencil_id, int P, int nfields, const T *const *in,
               T *const *out, size_t in_step, size_t out_step, cudaStream_t s,
               const FieldDev **dev_out) {
  constexpr unsigned EB = sizeof(T);
  constexpr bool f64 = EB == 8;
  FieldKey key;
  key.variant = stencil_id * 16 + P + (f64 ? 0 : 8);
  key.in_step = in_step;
  key.out_step = out_step;
  for (int f = 0; f < nfields; ++f) {
    key.ptrs.push_back(in[f]);
    key.ptrs.push_back(out[f]);
  }
  auto itf = plan->fields.find(key);
  if (itf != plan->fields.end()) {
    *dev_out = itf->second;
    return 0;
  }
  const int R = stencil_id == 13 || stencil_id == 125 ? 2 : (stencil_id == 25 ? 4 : 1);
  const bool cube = stencil_id == 125 || stencil_id == BRICKS_ST_27PT_COEFF;
  const int HW = (R > 2 || !f64) ? 4 : 2; /* i-face width in elements: a 16-byte multiple (StageLayout::HW) */
  const bool jswz = f64 && !cube && R == 4; /* swizzled four-row j-face box (march_dev.cuh StageLayout::JLX) */
  std::vector<FieldDev> host(nfields);
  for (int f = 0; f < nfields; ++f) {
    FieldDev &fd = host[f];
    memset(&fd, 0, sizeof fd);
    fd.out = (double *)out[f]; /* the kernel casts it back to its element type */
    fd.out_step = out_step;
    fd.in = (const double *)in[f];
    fd.in_step = in_step;
    for (int roll = 0; roll < 2; ++roll) {
      CUtensorMap *m = roll ? fd.maps.roll : fd.maps.main;
      const unsigned pk = roll ? R : P;
      /* the centre image is 64B-swizzled for fp64 (64-byte rows); fp32 rows are 32 bytes: dense */
      int rc = bricks::encode_map(&m[0], in[f], plan->nbricks, in_step, 8, 8, pk, f64, 1, EB);
      if (!rc) rc = bricks::encode_map(&m[1], in[f], plan->nbricks, in_step, HW, 8, pk, false, 1, EB);
      if (!rc) rc = bricks::encode_map(&m[2], in[f], plan->nbricks, in_step, 8, R, pk, false, 1, EB);
      if (!rc) rc = bricks::encode_map(&m[3], in[f], plan->nbricks, in_step, HW, cube ? R : 8, pk, false, 1, EB);
      if (!rc && jswz) rc = bricks::encode_map(&m[4], in[f], plan->nbricks, in_step, 8, 4, pk, true, 1, EB);
      if (rc) return rc;
    }
    int rc = bricks::encode_map(&fd.maps.pair[0], in[f], plan->nbricks, in_step, 8, 8, P, f64, 2, EB);
    if (!rc) rc = bricks::encode_map(&fd.maps.pair[1], in[f], plan->nbricks, in_step, 8, R, P, false, 2, EB);
    if (!rc && jswz) rc = bricks::encode_map(&fd.maps.pair[2], in[f], plan->nbricks, in_step, 8, 4, P, true, 2, EB);
    if (rc) return rc;
  }
  FieldDev *d = nullptr;
  BRK_CUDA(cudaMalloc(&d, sizeof(FieldDev) * nfields));
  BRK_CUDA(cudaMemcpyAsync(d, host.data(), sizeof(FieldDev) * nfields, cudaMemcpyHostToDevice, s));
  BRK_CUDA(cudaStreamSynchronize(s)); /* `host` dies at return; first use only */
  plan->fields[key] = d;
  *dev_out = d;
  return 0;
}

}  // namespace

namespace {

/* the body of bricks_stencil / bricks_stencil_f32: T = element type of the bricks */
template <typename T>
int run_stencil(bricks_plan *plan, int stencil_id, int nfields, const T *const *in, T *const *out, size_t in_step,
                size_t out_step, const double *coeff_host, int stream) {
  BRK_ARG(plan && in && out && nfields > 0 && nfields < 65536);
  BRK_ARG(stream >= 0 && stream < bricks::kNumStreams);
  BRK_ARG(in_step >= 512 && in_step % 2 == 0 && out_step >= 512 && out_step % 4 == 0);
  for (int f = 0; f < nfields; ++f) {
    BRK_ARG(in[f] && out[f] && in[f] != out[f]);
    /* TMA needs the field 16-byte aligned; a lane stores its quad as one 32-byte (fp64) / 16-byte (fp32) vector */
    BRK_ARG(((uintptr_t)in[f] & 15) == 0 && ((uintptr_t)out[f] & (4 * sizeof(T) - 1)) == 0);
  }
  switch (stencil_id) { /* before anything is built or cached for it */
    case BRICKS_ST_7PT: case BRICKS_ST_13PT: case BRICKS_ST_25PT: case BRICKS_ST_125PT: break;
    case BRICKS_ST_7PT_COEFF: case BRICKS_ST_COND: case BRICKS_ST_27PT_COEFF: BRK_ARG(coeff_host); break;
    default:
      bricks::set_error("unknown stencil id %d", stencil_id);
      return -1;
  }
  BRK_CUDA(cudaSetDevice(plan->ctx->device));
  cudaStream_t s = plan->ctx->streams[stream];
  constexpr int P = kPlanes;
  const FieldDev *fields = nullptr;
  int rc = get_fields(plan, stencil_id, P, nfields, in, out, in_step, out_step, s, &fields);
  if (rc) return rc;
  double c[27] = {0};
  switch (stencil_id) {
    case BRICKS_ST_7PT_COEFF:
      BRK_ARG(coeff_host);
      for (int i = 0; i < 7; ++i) c[i] = coeff_host[i];
      return launch<1, false, 1, P, T>(plan, stencil_id, fields, nfields, c, s);
    case BRICKS_ST_COND: /* cond.py (d3cond, stencils/3axis.cpp:94-163): coeff[0..6] as 7pt.py */
      BRK_ARG(coeff_host);
      for (int i = 0; i < 7; ++i) c[i] = coeff_host[i];
      return launch<1, false, 2, P, T>(plan, stencil_id, fields, nfields, c, s);
    case BRICKS_ST_27PT_COEFF: /* d3pt27 (stencils/3axis.cpp:165-247) */
      BRK_ARG(coeff_host);
      for (int i = 0; i < 27; ++i) c[i] = coeff_host[i];
      /* the reference reads term 16 at (k+1, j+1, i), the place of term 25, not at (k, j+1, i) */
      c[25] += c[16];
      c[16] = 0.0;
      return launch<1, true, 1, P, T>(plan, stencil_id, fields, nfields, c, s);
    case BRICKS_ST_7PT: /* MPI_ALPHA, MPI_BETA (stencils/fake.h:11-12) */
      c[0] = 0.4, c[1] = 0.1;
      return launch<1, false, 0, P, T>(plan, stencil_id, fields, nfields, c, s);
    case BRICKS_ST_13PT: /* MPI_B0..B2 (fake.h:20-22) */
      c[0] = 0.4, c[1] = 0.07, c[2] = 0.03;
      return launch<2, false, 0, P, T>(plan, stencil_id, fields, nfields, c, s);
    case BRICKS_ST_25PT: /* MPI_A0..A4 (fake.h:14-18) */
      c[0] = 0.1, c[1] = 0.06, c[2] = 0.045, c[3] = 0.03, c[4] = 0.015;
      return launch<4, false, 0, P, T>(plan, stencil_id, fields, nfields, c, s);
    case BRICKS_ST_125PT: { /* MPI_C0..C9 (fake.h:24-33) */
      const double cc[10] = {0.1, 0.04, 0.03, 0.01, 0.006, 0.004, 0.005, 0.002, 0.003, 0.001};
      for (int i = 0; i < 10; ++i) c[i] = cc[i];
      return launch<2, true, 0, P, T>(plan, stencil_id, fields, nfields, c, s);
    }
    default:
      bricks::set_error("unknown stencil id %d", stencil_id);
      return -1;
  }
}

}  // namespace

extern "C" int bricks_stencil(bricks_plan *plan, int stencil_id, int nfields, const double *const *in,
                              double *const *out, size_t in_step, size_t out_step,
                              const double *coeff_host, int stream) {
  return run_stencil<double>(plan, stencil_id, nfields, in, out, in_step, out_step, coeff_host, stream);
}

extern "C" int bricks_stencil_f32(bricks_plan *plan, int stencil_id, int nfields, const float *const *in,
                                  float *const *out, size_t in_step, size_t out_step,
                                  const double *coeff_host, int stream) {
  return run_stencil<float>(plan, stencil_id, nfields, in, out, in_step, out_step, coeff_host, stream);
}
