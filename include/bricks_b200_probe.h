/*
 * bricks_b200_probe.h -- hardware probes exported by libbricks_b200.so.
 * Not part of the drop-in boundary: these measure the roofs the stencil kernels are judged
 * against (fp64 FMA rate, HBM / L2 bandwidth of a plain copy, shared-memory vs shuffle
 * throughput) and check the TMA swizzle pattern the kernels assume.  bench.py and
 * profiles/ cite their output.
 */
#ifndef BRICKS_B200_PROBE_H
#define BRICKS_B200_PROBE_H
#include "bricks_b200.h"
#ifdef __cplusplus
extern "C" {
#endif
/* dependent-chain-free DFMA loop on every SM; returns achieved TFLOP/s (2 flops per FMA) */
int bricks_probe_dfma(bricks_ctx *ctx, int iters, double *tflops, float *ms);
/* out[i] = in[i] over `bytes` bytes, `reps` back-to-back launches; GB/s counts read + write */
int bricks_probe_copy(bricks_ctx *ctx, size_t bytes, int reps, double *gbs);
/* sum-reduce read of `bytes` bytes (no write traffic); GB/s counts reads */
int bricks_probe_read(bricks_ctx *ctx, size_t bytes, int reps, double *gbs);
/* mode 0: LDS.128 only, 1: SHFL only, 2: both interleaved; returns warp-instructions per
 * clock per SM for each of the two instruction kinds */
int bricks_probe_mio(bricks_ctx *ctx, int mode, double *lds_per_clk, double *shfl_per_clk);
/* TMA-load brick `b` of a field (step elements apart) with the 64B-swizzled centre map and
 * copy the raw 4096-byte shared-memory image to raw512 (host, 512 doubles) */
int bricks_probe_tma_image(bricks_ctx *ctx, const double *field_dev, unsigned nbricks, size_t step,
                           unsigned b, double *raw512);
/* cycles one warp spends per TMA copy it issues (2 KiB boxes, `nops` copies, multiple of 24):
 * mode 0 one lane in a loop, 1 one lane with 8 copies per iteration unrolled, 2 twelve lanes each
 * issuing their own copy; total = until the last byte has landed */
int bricks_probe_tma_issue(bricks_ctx *ctx, const double *field_dev, unsigned nbricks, size_t step, int mode,
                           int nops, double *issue_cycles_per_op, double *total_cycles_per_op);
#ifdef __cplusplus
}
#endif
#endif
