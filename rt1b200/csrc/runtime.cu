// Library-wide status plumbing and device queries (C ABI: include/rt1b200.h).
#include <stdarg.h>
#include <string.h>

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
