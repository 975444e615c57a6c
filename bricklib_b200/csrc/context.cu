// Context, memory, copies, streams, timers of libbricks_b200.so.
// Replaces the gpu* macro layer of the reference (include/brick-cuda.h:12-24,
// include/brick-gpu.h:23-99, stencils/cudaarray.h:11-31).
#include <sched.h>

#include <cctype>
#include <cstring>

#include "common.cuh"

namespace bricks {
static thread_local char g_err[1024] = "";
std::atomic<unsigned long long> g_launches{0};

void set_error(const char *fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof g_err, fmt, ap);
  va_end(ap);
}
}  // namespace bricks

using namespace bricks;

extern "C" {

const char *bricks_last_error(void) { return g_err; }

int bricks_device_count(int *count) {
  BRK_ARG(count);
  BRK_CUDA(cudaGetDeviceCount(count));
  return 0;
}

static int ctx_init(bricks_ctx *c, const cudaDeviceProp &prop) {
  c->sm_count = prop.multiProcessorCount;
  /* the communication stream outranks the compute stream, so exchange kernels (NCCL, region
   * copies) get the next free SM slots while a stencil sweep is in flight */
  int prio_lo = 0, prio_hi = 0;
  BRK_CUDA(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
  for (int s = 0; s < kNumStreams; ++s)
    BRK_CUDA(cudaStreamCreateWithPriority(&c->streams[s], cudaStreamNonBlocking,
                                          s == BRICKS_STREAM_COMM ? prio_hi : prio_lo));
  BRK_CUDA(cudaEventCreateWithFlags(&c->xev, cudaEventDisableTiming));
  for (int t = 0; t < kNumTimers; ++t) {
    BRK_CUDA(cudaEventCreate(&c->t0[t]));
    BRK_CUDA(cudaEventCreate(&c->t1[t]));
  }
  return 0;
}

int bricks_ctx_create(int device, bricks_ctx **out) {
  BRK_ARG(out);
  *out = nullptr;
  BRK_CUDA(cudaSetDevice(device));
  cudaDeviceProp prop;
  BRK_CUDA(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10) {
    set_error("bricks_b200 is built for sm_100a only; device %d is sm_%d%d", device, prop.major,
              prop.minor);
    return -2;
  }
  bricks_ctx *c = new bricks_ctx;
  c->device = device;
  const int rc = ctx_init(c, prop);
  if (rc) { /* keep the error text of the failing call; release what was created */
    bricks_ctx_destroy(c);
    return rc;
  }
  *out = c;
  return 0;
}

int bricks_ctx_destroy(bricks_ctx *c) {
  if (!c) return 0;
  cudaSetDevice(c->device);
  if (c->comm) ncclCommDestroy(c->comm);
  for (int s = 0; s < kNumStreams; ++s)
    if (c->streams[s]) cudaStreamDestroy(c->streams[s]);
  if (c->xev) cudaEventDestroy(c->xev);
  for (int t = 0; t < kNumTimers; ++t) {
    if (c->t0[t]) cudaEventDestroy(c->t0[t]);
    if (c->t1[t]) cudaEventDestroy(c->t1[t]);
  }
  delete c;
  return 0;
}

int bricks_ctx_device(const bricks_ctx *c, int *device) {
  BRK_ARG(c && device);
  *device = c->device;
  return 0;
}

static inline int pick_stream(bricks_ctx *c, int stream, cudaStream_t *s) {
  BRK_ARG(c && stream >= 0 && stream < kNumStreams);
  *s = c->streams[stream];
  return 0;
}

int bricks_malloc(bricks_ctx *c, void **dptr, size_t bytes) {
  BRK_ARG(c && dptr);
  BRK_CUDA(cudaSetDevice(c->device));
  BRK_CUDA(cudaMalloc(dptr, bytes ? bytes : 16));
  return 0;
}

int bricks_free(bricks_ctx *c, void *dptr) {
  BRK_ARG(c);
  BRK_CUDA(cudaSetDevice(c->device));
  BRK_CUDA(cudaFree(dptr));
  return 0;
}

int bricks_memset(bricks_ctx *c, void *dptr, int value, size_t bytes) {
  BRK_ARG(c && dptr);
  BRK_CUDA(cudaSetDevice(c->device));
  BRK_CUDA(cudaMemsetAsync(dptr, value, bytes, c->streams[BRICKS_STREAM_COMPUTE]));
  return 0;
}

int bricks_host_alloc(void **hptr, size_t bytes) {
  BRK_ARG(hptr);
  BRK_CUDA(cudaMallocHost(hptr, bytes ? bytes : 16));
  return 0;
}

/* CPUs of the NUMA node the device hangs off (sysfs), empty when the box does not say */
static bool device_local_cpus(int device, cpu_set_t *set, int *node_out) {
  char bus[32] = "";
  if (cudaDeviceGetPCIBusId(bus, sizeof bus, device) != cudaSuccess) return false;
  for (char *c = bus; *c; ++c) *c = (char)tolower(*c);
  char path[160];
  snprintf(path, sizeof path, "/sys/bus/pci/devices/%s/numa_node", bus);
  FILE *f = fopen(path, "r");
  int node = -1;
  if (f) {
    if (fscanf(f, "%d", &node) != 1) node = -1;
    fclose(f);
  }
  if (node < 0) return false;
  snprintf(path, sizeof path, "/sys/devices/system/node/node%d/cpulist", node);
  f = fopen(path, "r");
  if (!f) return false;
  char list[4096] = "";
  const bool got = fgets(list, sizeof list, f) != nullptr;
  fclose(f);
  if (!got) return false;
  CPU_ZERO(set);
  int n = 0;
  for (char *tok = strtok(list, ",\n"); tok; tok = strtok(nullptr, ",\n")) {
    int a = 0, b = 0;
    const int k = sscanf(tok, "%d-%d", &a, &b);
    if (k == 1) b = a;
    if (k < 1) continue;
    for (int c = a; c <= b && c < CPU_SETSIZE; ++c) CPU_SET(c, set), ++n;
  }
  if (node_out) *node_out = node;
  return n > 0;
}

int bricks_host_alloc_local(bricks_ctx *c, void **hptr, size_t bytes, int *numa_node) {
  BRK_ARG(c && hptr);
  if (numa_node) *numa_node = -1;
  BRK_CUDA(cudaSetDevice(c->device));
  /* first touch decides where pinned pages live: allocate and touch them from a CPU of the
   * device's NUMA node, then give the calling thread its own affinity back */
  cpu_set_t local, saved;
  int node = -1;
  const bool have = device_local_cpus(c->device, &local, &node) &&
                    sched_getaffinity(0, sizeof saved, &saved) == 0 &&
                    sched_setaffinity(0, sizeof local, &local) == 0;
  const cudaError_t e = cudaMallocHost(hptr, bytes ? bytes : 16);
  if (e == cudaSuccess) memset(*hptr, 0, bytes ? bytes : 16);
  if (have) sched_setaffinity(0, sizeof saved, &saved);
  BRK_CUDA(e);
  if (numa_node && have) *numa_node = node;
  return 0;
}

int bricks_host_free(void *hptr) {
  BRK_CUDA(cudaFreeHost(hptr));
  return 0;
}

static int copy_async(bricks_ctx *c, void *dst, const void *src, size_t bytes, int stream,
                      cudaMemcpyKind kind) {
  cudaStream_t s;
  int rc = pick_stream(c, stream, &s);
  if (rc) return rc;
  BRK_ARG(dst && src);
  BRK_CUDA(cudaSetDevice(c->device));
  BRK_CUDA(cudaMemcpyAsync(dst, src, bytes, kind, s));
  return 0;
}

int bricks_h2d(bricks_ctx *c, void *dst, const void *src, size_t bytes, int stream) {
  return copy_async(c, dst, src, bytes, stream, cudaMemcpyHostToDevice);
}
int bricks_d2h(bricks_ctx *c, void *dst, const void *src, size_t bytes, int stream) {
  return copy_async(c, dst, src, bytes, stream, cudaMemcpyDeviceToHost);
}
int bricks_d2d(bricks_ctx *c, void *dst, const void *src, size_t bytes, int stream) {
  return copy_async(c, dst, src, bytes, stream, cudaMemcpyDeviceToDevice);
}

int bricks_sync(bricks_ctx *c, int stream) {
  BRK_ARG(c && stream < kNumStreams);
  BRK_CUDA(cudaSetDevice(c->device));
  if (stream < 0)
    BRK_CUDA(cudaDeviceSynchronize());
  else
    BRK_CUDA(cudaStreamSynchronize(c->streams[stream]));
  return 0;
}

int bricks_stream_wait(bricks_ctx *c, int stream, int other) {
  BRK_ARG(c && stream >= 0 && stream < kNumStreams && other >= 0 && other < kNumStreams);
  BRK_CUDA(cudaSetDevice(c->device));
  BRK_CUDA(cudaEventRecord(c->xev, c->streams[other]));
  BRK_CUDA(cudaStreamWaitEvent(c->streams[stream], c->xev, 0));
  return 0;
}

int bricks_stream_handle(bricks_ctx *c, int stream, void **cuda_stream) {
  BRK_ARG(c && cuda_stream && stream >= 0 && stream < kNumStreams);
  *cuda_stream = (void *)c->streams[stream];
  return 0;
}

int bricks_timer_start(bricks_ctx *c, int stream, int slot) {
  BRK_ARG(c && slot >= 0 && slot < kNumTimers && stream >= 0 && stream < kNumStreams);
  BRK_CUDA(cudaEventRecord(c->t0[slot], c->streams[stream]));
  return 0;
}

int bricks_timer_stop(bricks_ctx *c, int stream, int slot) {
  BRK_ARG(c && slot >= 0 && slot < kNumTimers && stream >= 0 && stream < kNumStreams);
  BRK_CUDA(cudaEventRecord(c->t1[slot], c->streams[stream]));
  return 0;
}

int bricks_timer_elapsed_ms(bricks_ctx *c, int slot, float *ms) {
  BRK_ARG(c && ms && slot >= 0 && slot < kNumTimers);
  BRK_CUDA(cudaEventSynchronize(c->t1[slot]));
  BRK_CUDA(cudaEventElapsedTime(ms, c->t0[slot], c->t1[slot]));
  return 0;
}

int bricks_launch_count(unsigned long long *count) {
  BRK_ARG(count);
  *count = g_launches.load();
  return 0;
}

}  // extern "C"
