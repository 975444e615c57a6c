// Hardware probes (see include/bricks_b200_probe.h).  They establish the roofs the stencil
// kernels are measured against on the actual B200 and verify the TMA swizzle pattern that
// march_dev.cuh assumes.
#include "../../include/bricks_b200_probe.h"
#include "march_dev.cuh"

namespace {

__global__ void __launch_bounds__(256) dfma_kernel(double *out, int iters, double a, double b) {
  double r[8];
#pragma unroll
  for (int u = 0; u < 8; ++u) r[u] = threadIdx.x * 1e-9 + u;
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int u = 0; u < 8; ++u) r[u] = fma(r[u], a, b);
  }
  double s = 0;
#pragma unroll
  for (int u = 0; u < 8; ++u) s += r[u];
  if (s == 12345.678) out[0] = s;
}

__global__ void __launch_bounds__(256) copy_kernel(const int4 *__restrict__ in, int4 *__restrict__ out,
                                                   size_t n) {
  const size_t stride = (size_t)gridDim.x * blockDim.x * 4;
  for (size_t i = (size_t)blockIdx.x * blockDim.x * 4 + threadIdx.x; i < n; i += stride) {
    int4 v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (i + u * 256 < n) v[u] = in[i + u * 256];
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (i + u * 256 < n) out[i + u * 256] = v[u];
  }
}

__global__ void __launch_bounds__(256) read_kernel(const int4 *__restrict__ in, int *sink, size_t n) {
  const size_t stride = (size_t)gridDim.x * blockDim.x * 4;
  int acc = 0;
  for (size_t i = (size_t)blockIdx.x * blockDim.x * 4 + threadIdx.x; i < n; i += stride) {
    int4 v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) v[u] = (i + u * 256 < n) ? in[i + u * 256] : make_int4(0, 0, 0, 0);
#pragma unroll
    for (int u = 0; u < 4; ++u) acc ^= v[u].x ^ v[u].y ^ v[u].z ^ v[u].w;
  }
  if (acc == 0x12345677) sink[0] = acc;
}

/* mode bit0: LDS.128, bit1: SHFL.  8 warps per CTA, one CTA per SM. */
__global__ void __launch_bounds__(256) mio_kernel(int mode, int iters, long long *cycles, double *sink) {
  __shared__ double2 buf[1024];
  for (int i = threadIdx.x; i < 1024; i += 256) buf[i] = make_double2(i, -i);
  __syncthreads();
  double2 a = make_double2(0, 0);
  double s = threadIdx.x;
  const int lane = threadIdx.x & 31;
  const long long t0 = clock64();
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      if (mode & 1) {
        const double2 v = buf[(threadIdx.x + 32 * u + i) & 1023];
        a.x += v.x;
        a.y += v.y;
      }
      if (mode & 2) s += __shfl_sync(0xffffffffu, s, (lane + u + 1) & 31);
    }
  }
  const long long t1 = clock64();
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
  if (a.x + a.y + s == 1.2345e300) sink[0] = a.x;
}

__global__ void __launch_bounds__(32) tma_image_kernel(const __grid_constant__ CUtensorMap map, unsigned b,
                                                       double *out) {
  extern __shared__ __align__(1024) char smem[];
  const uint32_t s0 = dev::smem_u32(smem), bar = s0 + 4096;
  if (threadIdx.x == 0) {
    dev::mbar_init(bar, 1);
    dev::fence_barrier_init();
    dev::fence_proxy_async();
    dev::mbar_expect_tx(bar, 4096);
    dev::tma_load_4d(s0, &map, 0, 0, 0, (int)b, bar);
  }
  __syncwarp();
  dev::mbar_wait(bar, 0);
  for (int i = threadIdx.x; i < 512; i += 32) out[i] = reinterpret_cast<const double *>(smem)[i];
}


/* how many cycles one warp spends issuing TMA copies: mode 0 = one elected lane in a loop,
 * mode 1 = one elected lane, eight copies per iteration in straight-line code, mode 2 = twelve
 * lanes each issuing their own copy (the compiler's divergent-lane loop, as in march_kernel) */
__global__ void tma_issue_kernel(const __grid_constant__ CUtensorMap map, int mode, int nops, unsigned nbricks,
                                 long long *cycles) {
  extern __shared__ __align__(1024) char smem[];
  const uint32_t smem0 = dev::smem_u32(smem), bar = smem0 + 16 * 2048;
  const int lane = threadIdx.x;
  if (lane == 0) {
    dev::mbar_init(bar, 1);
    dev::fence_barrier_init();
    dev::fence_proxy_async();
    dev::mbar_expect_tx(bar, (uint32_t)nops * 2048u);
  }
  __syncwarp();
  const long long t0 = clock64();
  if (mode == 0) {
    if (lane == 0)
      for (int n = 0; n < nops; ++n)
        dev::tma_load_4d(smem0 + (n & 15) * 2048, &map, 0, 0, 4 * (n & 1), (int)((unsigned)(n * 7 + 1) % nbricks), bar);
  } else if (mode == 1) {
    if (lane == 0)
      for (int n = 0; n < nops; n += 8) {
#pragma unroll
        for (int u = 0; u < 8; ++u)
          dev::tma_load_4d(smem0 + ((n + u) & 15) * 2048, &map, 0, 0, 4 * (u & 1), (int)((unsigned)((n + u) * 7 + 1) % nbricks), bar);
      }
  } else {
    for (int n = 0; n < nops; n += 12) {
      if (lane < 12 && n + lane < nops)
        dev::tma_load_4d(smem0 + ((n + lane) & 15) * 2048, &map, 0, 0, 4 * (lane & 1), (int)((unsigned)((n + lane) * 7 + 1) % nbricks), bar);
      __syncwarp();
    }
  }
  __syncwarp();
  const long long t1 = clock64();
  dev::mbar_wait(bar, 0);
  const long long t2 = clock64();
  if (lane == 0) {
    cycles[0] = t1 - t0;
    cycles[1] = t2 - t0;
  }
}

}  // namespace

extern "C" {

int bricks_probe_dfma(bricks_ctx *ctx, int iters, double *tflops, float *ms_out) {
  BRK_ARG(ctx && tflops && iters > 0);
  BRK_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t s = ctx->streams[0];
  double *sink;
  BRK_CUDA(cudaMalloc(&sink, 64));
  const int ctas = ctx->sm_count * 8;
  dfma_kernel<<<ctas, 256, 0, s>>>(sink, 64, 1.0000001, 1e-9);
  BRK_CUDA(cudaEventRecord(ctx->t0[15], s));
  dfma_kernel<<<ctas, 256, 0, s>>>(sink, iters, 1.0000001, 1e-9);
  BRK_CUDA(cudaEventRecord(ctx->t1[15], s));
  BRK_CUDA(cudaEventSynchronize(ctx->t1[15]));
  float ms;
  BRK_CUDA(cudaEventElapsedTime(&ms, ctx->t0[15], ctx->t1[15]));
  *tflops = 2.0 * 8 * (double)iters * 256.0 * ctas / (ms * 1e-3) * 1e-12;
  if (ms_out) *ms_out = ms;
  bricks::g_launches += 2;
  BRK_CUDA(cudaFree(sink));
  return 0;
}

static int bw_probe(bricks_ctx *ctx, size_t bytes, int reps, double *gbs, bool write) {
  BRK_ARG(ctx && gbs && bytes >= 4096 && reps > 0);
  BRK_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t s = ctx->streams[0];
  char *a, *b;
  BRK_CUDA(cudaMalloc(&a, bytes));
  BRK_CUDA(cudaMalloc(&b, write ? bytes : 64));
  BRK_CUDA(cudaMemsetAsync(a, 1, bytes, s));
  const size_t n = bytes / 16;
  const int ctas = ctx->sm_count * 16;
  for (int r = 0; r < 3 + reps; ++r) {
    if (r == 3) BRK_CUDA(cudaEventRecord(ctx->t0[15], s));
    if (write)
      copy_kernel<<<ctas, 256, 0, s>>>((const int4 *)a, (int4 *)b, n);
    else
      read_kernel<<<ctas, 256, 0, s>>>((const int4 *)a, (int *)b, n);
  }
  BRK_CUDA(cudaEventRecord(ctx->t1[15], s));
  BRK_CUDA(cudaEventSynchronize(ctx->t1[15]));
  float ms;
  BRK_CUDA(cudaEventElapsedTime(&ms, ctx->t0[15], ctx->t1[15]));
  *gbs = (write ? 2.0 : 1.0) * (double)bytes * reps / (ms * 1e-3) * 1e-9;
  bricks::g_launches += 3 + reps;
  BRK_CUDA(cudaFree(a));
  BRK_CUDA(cudaFree(b));
  return 0;
}

int bricks_probe_copy(bricks_ctx *ctx, size_t bytes, int reps, double *gbs) {
  return bw_probe(ctx, bytes, reps, gbs, true);
}
int bricks_probe_read(bricks_ctx *ctx, size_t bytes, int reps, double *gbs) {
  return bw_probe(ctx, bytes, reps, gbs, false);
}

int bricks_probe_mio(bricks_ctx *ctx, int mode, double *lds_per_clk, double *shfl_per_clk) {
  BRK_ARG(ctx && lds_per_clk && shfl_per_clk && mode >= 1 && mode <= 3);
  BRK_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t s = ctx->streams[0];
  const int ctas = ctx->sm_count, iters = 4096;
  long long *cyc;
  double *sink;
  BRK_CUDA(cudaMalloc(&cyc, sizeof(long long) * ctas));
  BRK_CUDA(cudaMalloc(&sink, 64));
  mio_kernel<<<ctas, 256, 0, s>>>(mode, 64, cyc, sink);
  mio_kernel<<<ctas, 256, 0, s>>>(mode, iters, cyc, sink);
  std::vector<long long> h(ctas);
  BRK_CUDA(cudaMemcpyAsync(h.data(), cyc, sizeof(long long) * ctas, cudaMemcpyDeviceToHost, s));
  BRK_CUDA(cudaStreamSynchronize(s));
  double mean = 0;
  for (auto v : h) mean += (double)v;
  mean /= ctas;
  const double winst = 8.0 * iters * 8; /* warps * iters * unroll, per kind */
  *lds_per_clk = (mode & 1) ? winst / mean : 0.0;
  *shfl_per_clk = (mode & 2) ? winst / mean : 0.0;
  bricks::g_launches += 2;
  BRK_CUDA(cudaFree(cyc));
  BRK_CUDA(cudaFree(sink));
  return 0;
}

int bricks_probe_tma_image(bricks_ctx *ctx, const double *field_dev, unsigned nbricks, size_t step,
                           unsigned b, double *raw512) {
  BRK_ARG(ctx && field_dev && raw512 && b < nbricks);
  BRK_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t s = ctx->streams[0];
  CUtensorMap map;
  int rc = bricks::encode_map(&map, field_dev, nbricks, step, 8, 8, 8, true);
  if (rc) return rc;
  double *out;
  BRK_CUDA(cudaMalloc(&out, 4096));
  tma_image_kernel<<<1, 32, 4096 + 64, s>>>(map, b, out);
  BRK_CUDA(cudaGetLastError());
  BRK_CUDA(cudaMemcpyAsync(raw512, out, 4096, cudaMemcpyDeviceToHost, s));
  BRK_CUDA(cudaStreamSynchronize(s));
  ++bricks::g_launches;
  BRK_CUDA(cudaFree(out));
  return 0;
}

int bricks_probe_tma_issue(bricks_ctx *ctx, const double *field_dev, unsigned nbricks, size_t step, int mode,
                           int nops, double *issue_cycles_per_op, double *total_cycles_per_op) {
  BRK_ARG(ctx && field_dev && issue_cycles_per_op && total_cycles_per_op && mode >= 0 && mode <= 2);
  BRK_ARG(nops > 0 && nops % 24 == 0 && nops <= 480 && nbricks > 8);
  BRK_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t s = ctx->streams[0];
  CUtensorMap map;
  int rc = bricks::encode_map(&map, field_dev, nbricks, step, 8, 8, 4, true);
  if (rc) return rc;
  long long *cyc;
  BRK_CUDA(cudaMalloc(&cyc, 16));
  const int smem = 16 * 2048 + 64;
  BRK_CUDA(cudaFuncSetAttribute(tma_issue_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  for (int rep = 0; rep < 2; ++rep) tma_issue_kernel<<<1, 32, smem, s>>>(map, mode, nops, nbricks, cyc);
  BRK_CUDA(cudaGetLastError());
  long long h[2];
  BRK_CUDA(cudaMemcpyAsync(h, cyc, 16, cudaMemcpyDeviceToHost, s));
  BRK_CUDA(cudaStreamSynchronize(s));
  bricks::g_launches += 2;
  *issue_cycles_per_op = (double)h[0] / nops;
  *total_cycles_per_op = (double)h[1] / nops;
  BRK_CUDA(cudaFree(cyc));
  return 0;
}

}  // extern "C"
