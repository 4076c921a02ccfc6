// rt1b200 — shared device/host helpers for the sm_100a kernels.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#define RT1_API extern "C" __attribute__((visibility("default")))

// ---- host-side status plumbing --------------------------------------------------------------------------
void rt1_set_error(const char* fmt, ...);
int rt1_check_launch(const char* what);   // cudaGetLastError() → status + message
int rt1_ensure_dyn_smem(const void* kernel, size_t bytes);   // per-(kernel, device) opt-in to > 48 KB dynamic shared memory
int rt1_sm_count();                       // SM count of the current device (cached per device)
// Device word added to every dropout seed (or null): lets a CAPTURED step draw fresh masks on every replay — the seeds
// baked into the graph's kernel arguments stay fixed, the word is advanced by a node of the graph itself.
const uint32_t* rt1_seed_epoch_ptr();

#define RT1_CHECK_CUDA(expr)                                                                  \
  do {                                                                                        \
    cudaError_t _e = (expr);                                                                  \
    if (_e != cudaSuccess) {                                                                  \
      rt1_set_error("%s failed at %s:%d: %s", #expr, __FILE__, __LINE__, cudaGetErrorString(_e)); \
      return (int)_e;                                                                         \
    }                                                                                         \
  } while (0)

#define RT1_REQUIRE(cond, ...)            \
  do {                                    \
    if (!(cond)) {                        \
      rt1_set_error(__VA_ARGS__);         \
      return -1;                          \
    }                                     \
  } while (0)

typedef __nv_bfloat16 bf16;
typedef __nv_bfloat162 bf162;

static inline int rt1_cdiv(long a, long b) { return (int)((a + b - 1) / b); }

// ---- programmatic dependent launch (opt-in: RT1B200_PDL=1) -------------------------------------------------
// A training step is ~900 kernels of 5-50 us in ONE dependency chain.  With programmatic stream serialization a kernel's
// CTAs may be placed while the previous kernel of the stream is still draining; they run whatever does not touch global
// memory and block in pdl_wait() until the previous grid has completed and its writes are visible.  pdl_enter() therefore
// precedes EVERY global access of every kernel launched through rt1_launch (reads, and writes too: the allocator may hand
// a buffer the predecessor still reads to its successor as output); in the tcgen05 kernels it comes after the TMEM
// allocation, so that a placed-and-waiting CTA can never hold columns a running kernel still has to allocate.
// Measured on B200 (profiles/r2_pdl_modes.jsonl): 18.67-18.69 ms per captured step with it against 18.81 without, and
// 9.0 vs 9.2 us per kernel on a chain of 400 small dependent kernels — within noise, so it is OFF by default; triggering
// before waiting (mode 3: arbitrarily deep chains of placed-and-waiting grids) doubled the chain's time.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
// What a kernel does where its prologue ends (build-time choice RT1B200_PDL_MODE, see rt1b200/build.py):
//   1  wait, then trigger: the next kernel's CTAs are placed while this one runs its main loop (one kernel deep)
//   2  wait only: the trigger is the implicit one at CTA exit — the next kernel's CTAs fill in as this one's drain
//   3  trigger, then wait
#ifndef RT1_PDL_MODE
#define RT1_PDL_MODE 2
#endif
__device__ __forceinline__ void pdl_enter() {
#if RT1_PDL_MODE == 1
  pdl_wait();
  pdl_trigger();
#elif RT1_PDL_MODE == 2
  pdl_wait();
#elif RT1_PDL_MODE == 3
  pdl_trigger();
  pdl_wait();
#endif
}
bool rt1_pdl_enabled();

template <typename... KArgs, typename... Args>
static inline void rt1_launch(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = rt1_pdl_enabled() ? 1 : 0;
  (void)cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);   // the status is collected by rt1_check_launch()
}

// ---- device helpers -------------------------------------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// Block-wide sum for blockDim.x <= 1024; `scratch` must hold 32 floats. All threads get the result.
__device__ __forceinline__ float block_sum(float v, float* scratch) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  v = warp_sum(v);
  __syncthreads();  // protect scratch reuse between consecutive calls
  if (lane == 0) scratch[wid] = v;
  __syncthreads();
  float r = (lane < nw) ? scratch[lane] : 0.f;
  r = warp_sum(r);
  return r;
}
__device__ __forceinline__ float block_max(float v, float* scratch) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  v = warp_max(v);
  __syncthreads();
  if (lane == 0) scratch[wid] = v;
  __syncthreads();
  float r = (lane < nw) ? scratch[lane] : -INFINITY;
  r = warp_max(r);
  return r;
}

// Stateless dropout RNG: lowbias32 avalanche hash of (index + seed * golden). Mirrors
// rt1b200/ops/reference.py::hash_u32 bit for bit so tests can rebuild the exact mask on the host.
__device__ __forceinline__ uint32_t rt1_hash_u32(uint32_t idx, uint32_t seed) {
  uint32_t x = idx + seed * 0x9E3779B9u;
  x ^= x >> 16;
  x *= 0x7FEB352Du;
  x ^= x >> 15;
  x *= 0x846CA68Bu;
  x ^= x >> 16;
  return x;
}
__host__ __device__ __forceinline__ uint32_t rt1_drop_threshold(float p) {
  double t = (double)p * 4294967296.0;
  return t >= 4294967295.0 ? 0xFFFFFFFFu : (uint32_t)t;
}

// sigmoid(x) = 0.5 * tanh(x / 2) + 0.5 with the hardware tanh: ONE MUFU op (MUFU.TANH) instead of MUFU.EX2 + MUFU.RCP.
// The SFU issues 16 lanes/clk/SM against 128 for FMA, so in the elementwise BN+SiLU passes every MUFU op costs as much
// issue bandwidth as 8 FMAs.  tanh.approx.f32 has a relative error of ~2^-11: below bf16 output rounding (2^-9).
__device__ __forceinline__ float tanh_approx_f(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float sigmoid_f(float x) { return fmaf(tanh_approx_f(0.5f * x), 0.5f, 0.5f); }
__device__ __forceinline__ float silu_f(float x) { return x * sigmoid_f(x); }
// d/dx silu(x) = s * (1 + x * (1 - s))
__device__ __forceinline__ float silu_grad_f(float x) {
  float s = sigmoid_f(x);
  return s * (1.f + x * (1.f - s));
}
__device__ __forceinline__ float gelu_tanh_f(float x) {
  const float k0 = 0.7978845608028654f, k1 = 0.044715f;
  float u = k0 * (x + k1 * x * x * x);
  return 0.5f * x * (1.f + tanhf(u));
}
__device__ __forceinline__ float gelu_tanh_grad_f(float x) {
  const float k0 = 0.7978845608028654f, k1 = 0.044715f;
  float u = k0 * (x + k1 * x * x * x);
  float t = tanhf(u);
  float du = k0 * (1.f + 3.f * k1 * x * x);
  return 0.5f * (1.f + t) + 0.5f * x * (1.f - t * t) * du;
}

// 16-byte vector of 8 bf16
struct __align__(16) bf16x8 {
  bf162 v[4];
};
__device__ __forceinline__ void unpack8(const bf16x8& p, float* f) {
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float2 t = __bfloat1622float2(p.v[i]);
    f[2 * i] = t.x;
    f[2 * i + 1] = t.y;
  }
}
__device__ __forceinline__ bf16x8 pack8(const float* f) {
  bf16x8 p;
#pragma unroll
  for (int i = 0; i < 4; ++i) p.v[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
  return p;
}
