// Library-wide status plumbing and device queries (C ABI: include/rt1b200.h).
#include <stdarg.h>
#include <string.h>

#include <mutex>

#include "common.cuh"

static thread_local char g_err[1024] = "";

void rt1_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int rt1_check_launch(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    rt1_set_error("%s: %s", what, cudaGetErrorString(e));
    return (int)e;
  }
  return 0;
}

// cudaFuncSetAttribute and device attributes are PER DEVICE: caches are keyed by the current device so that a second GPU
// driven from the same process gets its own opt-in to > 48 KB of dynamic shared memory.
namespace {
constexpr int kMaxDev = 64, kMaxKernels = 256;
struct SmemOptIn {
  const void* kernel;
  size_t bytes[kMaxDev];
};
SmemOptIn g_optin[kMaxKernels];
int g_optin_n = 0;
std::mutex g_optin_mu;
int g_sm_count[kMaxDev] = {0};
}  // namespace

int rt1_ensure_dyn_smem(const void* kernel, size_t bytes) {
  if (bytes <= 48 * 1024) return 0;
  int dev = 0;
  RT1_CHECK_CUDA(cudaGetDevice(&dev));
  std::lock_guard<std::mutex> lock(g_optin_mu);
  SmemOptIn* e = nullptr;
  for (int i = 0; i < g_optin_n; ++i)
    if (g_optin[i].kernel == kernel) e = &g_optin[i];
  if (e == nullptr && g_optin_n < kMaxKernels) {
    e = &g_optin[g_optin_n++];
    e->kernel = kernel;
    memset(e->bytes, 0, sizeof(e->bytes));
  }
  if (e != nullptr && dev >= 0 && dev < kMaxDev && e->bytes[dev] >= bytes) return 0;
  RT1_CHECK_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
  if (e != nullptr && dev >= 0 && dev < kMaxDev) e->bytes[dev] = bytes;
  return 0;
}

bool rt1_pdl_enabled() {
  static int on = -1;
  if (on < 0) {
    const char* e = getenv("RT1B200_PDL");
    on = (e != nullptr && e[0] == '1') ? 1 : 0;     // opt-in: measured neutral on the captured step (profiles/ncu_notes.md §12)
  }
  return on == 1;
}

int rt1_sm_count() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDev) return 148;
  if (g_sm_count[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    g_sm_count[dev] = n;
  }
  return g_sm_count[dev];
}

namespace {
const uint32_t* g_seed_epoch[kMaxDev] = {nullptr};
}
const uint32_t* rt1_seed_epoch_ptr() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDev) return nullptr;
  return g_seed_epoch[dev];
}
// Register (or clear, with null) the current device's seed-epoch word: a caller-owned uint32 in device memory.
RT1_API int rt1_set_seed_epoch(const void* dev_u32) {
  int dev = 0;
  RT1_CHECK_CUDA(cudaGetDevice(&dev));
  RT1_REQUIRE(dev >= 0 && dev < kMaxDev, "rt1_set_seed_epoch: device %d out of range", dev);
  g_seed_epoch[dev] = (const uint32_t*)dev_u32;
  return 0;
}

RT1_API const char* rt1_last_error(void) { return g_err; }
RT1_API int rt1_abi_version(void) { return 1; }

RT1_API int rt1_device_info(int* sm_count, int* cc_major, int* cc_minor, long long* smem_optin) {
  int dev = 0;
  RT1_CHECK_CUDA(cudaGetDevice(&dev));
  cudaDeviceProp p;
  RT1_CHECK_CUDA(cudaGetDeviceProperties(&p, dev));
  if (sm_count) *sm_count = p.multiProcessorCount;
  if (cc_major) *cc_major = p.major;
  if (cc_minor) *cc_minor = p.minor;
  if (smem_optin) *smem_optin = (long long)p.sharedMemPerBlockOptin;
  return 0;
}

// L2 flush helper for benchmarks: streams `bytes` of writes through a caller-provided buffer.
__global__ void l2_flush_kernel(uint4* buf, long n16, uint32_t tag) {
  long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  long stride = (long)gridDim.x * blockDim.x;
  for (; i < n16; i += stride) buf[i] = make_uint4(tag, tag, tag, tag);
}
RT1_API int rt1_l2_flush(void* buf, long long bytes, unsigned tag, cudaStream_t stream) {
  long n16 = bytes / 16;
  if (n16 <= 0) return 0;
  l2_flush_kernel<<<148 * 8, 256, 0, stream>>>((uint4*)buf, n16, tag);
  return rt1_check_launch("rt1_l2_flush");
}

// Publish a device scalar to the host from INSIDE a captured graph: the value goes to pinned (device-mapped) host memory,
// then a step counter is bumped and written next to it.  The host polls the counter: it learns the step's loss as soon as
// the forward pass has produced it, while the rest of the graph (backward, optimizer) is still running — an event cannot be
// waited on in the middle of a graph launch.
__global__ void publish_scalar_kernel(const float* __restrict__ src, float* dst_host, unsigned* counter, volatile unsigned* flag_host) {
  pdl_enter();
  *dst_host = *src;
  __threadfence_system();
  const unsigned c = *counter + 1u;
  *counter = c;
  *flag_host = c;
}
RT1_API int rt1_publish_scalar(const float* src_dev, float* dst_host, unsigned* counter_dev, unsigned* flag_host, cudaStream_t stream) {
  RT1_REQUIRE(src_dev != nullptr && dst_host != nullptr && counter_dev != nullptr && flag_host != nullptr, "rt1_publish_scalar: null pointer");
  publish_scalar_kernel<<<1, 1, 0, stream>>>(src_dev, dst_host, counter_dev, flag_host);
  return rt1_check_launch("rt1_publish_scalar");
}
