// NVLink/NVSwitch peer-to-peer collectives fused with the optimizer (SURVEY §8 A1–A3).
//
// Replaces, on the one communication-bound hot path of RT-1 training, the whole chain
//     DDP bucket copy → ncclAllReduce → divide by world → torch.optim.Adam (foreach kernels) → zero_grad
// (reference distribute_train.py:99-110,231-238) with ONE kernel per gradient bucket that talks to the peers'
// memory directly.  Every rank's bf16 gradient and bf16 parameter buffers live in "symmetric" allocations that
// all peers have mapped (cudaIpc handles exchanged once at start-up); B200's NVSwitch gives each GPU full
// bandwidth to every peer, so a two-shot schedule is bandwidth-optimal without any ring:
//
//   barrier "ready"   : flag write to every peer (st.release.sys) + wait on own flags (ld.acquire.sys)
//   reduce-scatter    : rank r PULLS shard r of the bucket from every peer (128-bit loads over NVLink), sums in fp32
//   optimizer         : x 1/world, bf16→fp32, AdamW on the fp32 master / moment shards that rank r owns
//   all-gather        : rank r PUSHES the updated bf16 parameter shard into every peer's parameter buffer
//   barrier "done"    : last CTA of the rank signals; afterwards each rank clears its own gradient bucket
//
// so per element a rank moves 2(W-1)/W · 2 B over NVLink and touches its fp32 state only for 1/W of the bucket
// (the sharded update SURVEY App. C argues for).  The same file provides all-reduce-only kernels (one-shot for
// latency-bound sizes, two-shot for bandwidth-bound ones) used for scalars and for the 1 KB–256 MB sweep.
//
// Test/emulation mode: with gridDim.y == world the kernel plays ALL ranks in one launch on one GPU (rank =
// blockIdx.y, all CTAs co-resident), which is how the multi-rank logic is verified on a single B200 — separate
// launches that spin on each other on one GPU are not safe.
#include <string.h>

#include "adam.cuh"

#define P2P_MAX_WORLD 8
#define P2P_THREADS 512

struct P2PComm {
  int world;
  void* grad[P2P_MAX_WORLD];        // bf16 gradient buffers (symmetric), indexed by rank
  void* param[P2P_MAX_WORLD];       // bf16 parameter buffers (symmetric)
  unsigned* flags[P2P_MAX_WORLD];   // signal pads: [2][P2P_MAX_WORLD] epochs (ready, done) + [1] CTA counter
};
struct P2PLocal {
  float* master[P2P_MAX_WORLD];     // fp32 state; only [rank] is valid outside emulation mode
  float* exp_avg[P2P_MAX_WORLD];
  float* exp_avg_sq[P2P_MAX_WORLD];
};

__device__ __forceinline__ void st_release_sys(unsigned* p, unsigned v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint4 ld_peer_v4(const void* p) {  // volatile: never served from a stale L1 line
  uint4 v;
  asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}

// flags layout per rank: [0 .. W) ready epochs, [MAX .. MAX+W) done epochs, [2*MAX] CTA arrival counter
__device__ __forceinline__ void signal_all(const P2PComm& c, int rank, int slot, unsigned epoch) {
  if ((int)threadIdx.x < c.world) st_release_sys(c.flags[threadIdx.x] + slot * P2P_MAX_WORLD + rank, epoch);
}
__device__ __forceinline__ void wait_all(const P2PComm& c, int rank, int slot, unsigned epoch) {
  if ((int)threadIdx.x < c.world) {
    const unsigned* f = c.flags[rank] + slot * P2P_MAX_WORLD + threadIdx.x;
    unsigned spins = 0;
    while ((int)(ld_acquire_sys(f) - epoch) < 0) {
      if (++spins > (1u << 28)) {
        printf("rt1b200 p2p: rank %d timed out waiting for peer %d (slot %d epoch %u, saw %u)\n", rank, threadIdx.x, slot, epoch, *f);
        __trap();
      }
    }
  }
  __syncthreads();
}
// grid-wide "everyone on this rank has finished" → the last CTA to arrive signals the peers
__device__ __forceinline__ void rank_done_signal(const P2PComm& c, int rank, unsigned epoch) {
  __shared__ int is_last;
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned* counter = c.flags[rank] + 2 * P2P_MAX_WORLD;
    const unsigned prev = atomicAdd(counter, 1u);
    is_last = (prev == gridDim.x - 1);
    if (is_last) *counter = 0;  // re-arm for the next launch (stream order protects it)
  }
  __syncthreads();
  if (is_last) signal_all(c, rank, 1, epoch);
}

__device__ __forceinline__ void shard_bounds(long n, int world, int r, long& lo, long& hi) {
  long chunk = ((n + world - 1) / world + 7) & ~7L;
  lo = min(n, chunk * r);
  hi = min(n, lo + chunk);
}

__global__ void __launch_bounds__(P2P_THREADS) p2p_allreduce_adamw_kernel(P2PComm c, P2PLocal loc, int base_rank, long off, long n,
                                                                         unsigned epoch, AdamHyper h) {
  const int rank = base_rank + blockIdx.y, W = c.world;
  if (blockIdx.x == 0) signal_all(c, rank, 0, epoch);
  wait_all(c, rank, 0, epoch);

  long lo, hi;
  shard_bounds(n, W, rank, lo, hi);
  float* master = loc.master[rank] + off;
  float* m = loc.exp_avg[rank] + off;
  float* v = loc.exp_avg_sq[rank] + off;
  const long nvec = (hi - lo) >> 3;
  const long tstride = (long)gridDim.x * blockDim.x;
  // Software-pipelined by one vector: the peers' loads for iteration i+1 are in flight (NVLink round trip
  // ~2 us) while iteration i does its fp32 optimizer math and pushes parameters.
  uint4 raw[P2P_MAX_WORLD], nxt[P2P_MAX_WORLD];
  long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < nvec) {
#pragma unroll
    for (int p = 0; p < P2P_MAX_WORLD; ++p)
      if (p < W) nxt[p] = ld_peer_v4(reinterpret_cast<const bf16*>(c.grad[(rank + p) % W]) + off + lo + (i << 3));
  }
  for (; i < nvec; i += tstride) {
    const long e = lo + (i << 3);
    float g[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
    for (int p = 0; p < P2P_MAX_WORLD; ++p)
      if (p < W) raw[p] = nxt[p];
    if (i + tstride < nvec) {
#pragma unroll
      for (int p = 0; p < P2P_MAX_WORLD; ++p)
        if (p < W) nxt[p] = ld_peer_v4(reinterpret_cast<const bf16*>(c.grad[(rank + p) % W]) + off + lo + ((i + tstride) << 3));
    }
#pragma unroll
    for (int p = 0; p < P2P_MAX_WORLD; ++p)
      if (p < W) {
        float f[8];
        unpack8(*reinterpret_cast<const bf16x8*>(&raw[p]), f);
#pragma unroll
        for (int k = 0; k < 8; ++k) g[k] += f[k];
      }
    float4 p0 = reinterpret_cast<float4*>(master + e)[0], p1 = reinterpret_cast<float4*>(master + e)[1];
    float4 m0 = reinterpret_cast<float4*>(m + e)[0], m1 = reinterpret_cast<float4*>(m + e)[1];
    float4 v0 = reinterpret_cast<float4*>(v + e)[0], v1 = reinterpret_cast<float4*>(v + e)[1];
    float pp[8] = {p0.x, p0.y, p0.z, p0.w, p1.x, p1.y, p1.z, p1.w};
    float mm[8] = {m0.x, m0.y, m0.z, m0.w, m1.x, m1.y, m1.z, m1.w};
    float vv[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
#pragma unroll
    for (int k = 0; k < 8; ++k) adam_update(pp[k], g[k], mm[k], vv[k], h);
    reinterpret_cast<float4*>(master + e)[0] = make_float4(pp[0], pp[1], pp[2], pp[3]);
    reinterpret_cast<float4*>(master + e)[1] = make_float4(pp[4], pp[5], pp[6], pp[7]);
    reinterpret_cast<float4*>(m + e)[0] = make_float4(mm[0], mm[1], mm[2], mm[3]);
    reinterpret_cast<float4*>(m + e)[1] = make_float4(mm[4], mm[5], mm[6], mm[7]);
    reinterpret_cast<float4*>(v + e)[0] = make_float4(vv[0], vv[1], vv[2], vv[3]);
    reinterpret_cast<float4*>(v + e)[1] = make_float4(vv[4], vv[5], vv[6], vv[7]);
    const bf16x8 packed = pack8(pp);
#pragma unroll
    for (int p = 0; p < P2P_MAX_WORLD; ++p)
      if (p < W) *reinterpret_cast<bf16x8*>(reinterpret_cast<bf16*>(c.param[(rank + p) % W]) + off + e) = packed;
  }
  rank_done_signal(c, rank, epoch);
  wait_all(c, rank, 1, epoch);
  // every peer has consumed this rank's gradients: clear the bucket for the next step
  uint4* gz = reinterpret_cast<uint4*>(reinterpret_cast<bf16*>(c.grad[rank]) + off);
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < (n >> 3); i += (long)gridDim.x * blockDim.x)
    gz[i] = make_uint4(0, 0, 0, 0);
}

// In-place sum all-reduce of a symmetric buffer (bf16 or fp32), scaled by `scale`.
//   two_shot = 1: reduce-scatter by pulling + all-gather by pushing (bandwidth-optimal on NVSwitch)
//   two_shot = 0: every rank pulls everything and writes only its own copy (one barrier less per byte; wins
//                 below a few hundred KB where the barriers dominate)
template <bool IS_BF16>
__global__ void __launch_bounds__(P2P_THREADS) p2p_allreduce_kernel(P2PComm c, int base_rank, long off_bytes, long n, unsigned epoch,
                                                                   float scale, int two_shot) {
  const int rank = base_rank + blockIdx.y, W = c.world;
  constexpr int EPV = IS_BF16 ? 8 : 4;  // elements per 16-byte vector
  if (blockIdx.x == 0) signal_all(c, rank, 0, epoch);
  wait_all(c, rank, 0, epoch);
  long lo = 0, hi = n;
  if (two_shot) {
    long chunk = ((n + W - 1) / W + EPV - 1) / EPV * EPV;
    lo = min(n, chunk * rank);
    hi = min(n, lo + chunk);
  }
  const long nvec = (hi - lo) / EPV;
  const size_t esz = IS_BF16 ? 2 : 4;
  // one-shot: every rank reads all of [0, n) from every peer, so nobody may overwrite the inputs while a peer
  // could still be reading them — the result goes to elements [n, 2n) of the rank's own buffer instead.
  const long tstride = (long)gridDim.x * blockDim.x;
  uint4 nxt[P2P_MAX_WORLD];
  long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < nvec) {
#pragma unroll
    for (int p = 0; p < P2P_MAX_WORLD; ++p)
      if (p < W) nxt[p] = ld_peer_v4(reinterpret_cast<const char*>(c.grad[(rank + p) % W]) + off_bytes + (lo + i * EPV) * esz);
  }
  for (; i < nvec; i += tstride) {
    const long e = lo + i * EPV;
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    uint4 cur[P2P_MAX_WORLD];
#pragma unroll
    for (int p = 0; p < P2P_MAX_WORLD; ++p)
      if (p < W) cur[p] = nxt[p];
    if (i + tstride < nvec) {   // prefetch the next vector from every peer before reducing this one
#pragma unroll
      for (int p = 0; p < P2P_MAX_WORLD; ++p)
        if (p < W) nxt[p] = ld_peer_v4(reinterpret_cast<const char*>(c.grad[(rank + p) % W]) + off_bytes + (e + tstride * EPV) * esz);
    }
#pragma unroll
    for (int p = 0; p < P2P_MAX_WORLD; ++p)
      if (p < W) {
        const uint4 raw = cur[p];
        if (IS_BF16) {
          float f[8];
          unpack8(*reinterpret_cast<const bf16x8*>(&raw), f);
#pragma unroll
          for (int k = 0; k < 8; ++k) acc[k] += f[k];
        } else {
          acc[0] += __uint_as_float(raw.x); acc[1] += __uint_as_float(raw.y);
          acc[2] += __uint_as_float(raw.z); acc[3] += __uint_as_float(raw.w);
        }
      }
#pragma unroll
    for (int k = 0; k < 8; ++k) acc[k] *= scale;
    uint4 outv;
    if (IS_BF16) {
      const bf16x8 pk = pack8(acc);
      outv = *reinterpret_cast<const uint4*>(&pk);
    } else {
      outv = make_uint4(__float_as_uint(acc[0]), __float_as_uint(acc[1]), __float_as_uint(acc[2]), __float_as_uint(acc[3]));
    }
    if (two_shot) {
#pragma unroll
      for (int p = 0; p < P2P_MAX_WORLD; ++p)
        if (p < W) *reinterpret_cast<uint4*>(reinterpret_cast<char*>(c.grad[(rank + p) % W]) + off_bytes + e * esz) = outv;
    } else {
      *reinterpret_cast<uint4*>(reinterpret_cast<char*>(c.grad[rank]) + off_bytes + (n + e) * esz) = outv;
    }
  }
  rank_done_signal(c, rank, epoch);
  wait_all(c, rank, 1, epoch);
}

// ---------------------------------------------------------------------------------------------- host side
RT1_API int rt1_symm_alloc(long long bytes, void** ptr, unsigned char* handle64) {
  RT1_CHECK_CUDA(cudaMalloc(ptr, (size_t)bytes));
  RT1_CHECK_CUDA(cudaMemset(*ptr, 0, (size_t)bytes));
  cudaIpcMemHandle_t h;
  RT1_CHECK_CUDA(cudaIpcGetMemHandle(&h, *ptr));
  memcpy(handle64, &h, sizeof(h));
  return 0;
}
RT1_API int rt1_symm_open(const unsigned char* handle64, void** ptr) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, sizeof(h));
  RT1_CHECK_CUDA(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return 0;
}
RT1_API int rt1_symm_close(void* ptr) {
  RT1_CHECK_CUDA(cudaIpcCloseMemHandle(ptr));
  return 0;
}
RT1_API int rt1_symm_free(void* ptr) {
  RT1_CHECK_CUDA(cudaFree(ptr));
  return 0;
}
RT1_API int rt1_p2p_flag_words(void) { return 2 * P2P_MAX_WORLD + 8; }

static int p2p_fill(P2PComm& c, int world, void* const* grad, void* const* param, void* const* flags) {
  RT1_REQUIRE(world >= 1 && world <= P2P_MAX_WORLD, "p2p: world size %d not in [1, %d]", world, P2P_MAX_WORLD);
  c.world = world;
  for (int i = 0; i < P2P_MAX_WORLD; ++i) {
    c.grad[i] = i < world ? grad[i] : nullptr;
    c.param[i] = (i < world && param) ? param[i] : nullptr;
    c.flags[i] = i < world ? (unsigned*)flags[i] : nullptr;
  }
  return 0;
}

// grad/param/flags: host arrays of `world` device pointers (peer-mapped).  master/exp_avg/exp_avg_sq: host arrays
// of device pointers; outside emulation only entry [rank] is used.  emulate_all_ranks = 1 plays every rank in one
// launch on the current device (single-GPU test mode; needs world * ctas_per_rank co-resident CTAs).
RT1_API int rt1_p2p_allreduce_adamw(int world, int rank, void* const* grad, void* const* param, void* const* flags,
                                    float* const* master, float* const* exp_avg, float* const* exp_avg_sq, long long offset,
                                    long long n, unsigned epoch, float lr, float beta1, float beta2, float eps, float weight_decay,
                                    float bias_corr1, float bias_corr2, int ctas_per_rank, int emulate_all_ranks,
                                    cudaStream_t stream) {
  P2PComm c;
  int rc = p2p_fill(c, world, grad, param, flags);
  if (rc) return rc;
  RT1_REQUIRE(n % 8 == 0 && offset % 8 == 0, "p2p_allreduce_adamw: offset/length must be multiples of 8 elements");
  P2PLocal loc;
  for (int i = 0; i < P2P_MAX_WORLD; ++i) {
    const bool valid = emulate_all_ranks ? i < world : i == rank;
    loc.master[i] = valid ? master[i] : nullptr;
    loc.exp_avg[i] = valid ? exp_avg[i] : nullptr;
    loc.exp_avg_sq[i] = valid ? exp_avg_sq[i] : nullptr;
  }
  AdamHyper h{lr, beta1, beta2, eps, weight_decay, 1.f / bias_corr1, 1.f / sqrtf(bias_corr2), 1.f / (float)world};
  if (ctas_per_rank < 1) ctas_per_rank = 32;
  dim3 grid(ctas_per_rank, emulate_all_ranks ? world : 1);
  p2p_allreduce_adamw_kernel<<<grid, P2P_THREADS, 0, stream>>>(c, loc, emulate_all_ranks ? 0 : rank, (long)offset, (long)n, epoch, h);
  return rt1_check_launch("rt1_p2p_allreduce_adamw");
}

// In-place all-reduce (sum * scale) of `n` elements at byte offset `offset_bytes` of the symmetric buffers `buf`.
// one-shot (two_shot = 0) writes the result to elements [n, 2n) of the rank's own buffer instead.
RT1_API int rt1_p2p_allreduce(int world, int rank, void* const* buf, void* const* flags, long long offset_bytes, long long n,
                              int is_bf16, float scale, int two_shot, unsigned epoch, int ctas_per_rank, int emulate_all_ranks,
                              cudaStream_t stream) {
  P2PComm c;
  int rc = p2p_fill(c, world, buf, nullptr, flags);
  if (rc) return rc;
  RT1_REQUIRE(n % (is_bf16 ? 8 : 4) == 0 && offset_bytes % 16 == 0, "p2p_allreduce: 16-byte granularity required");
  if (ctas_per_rank < 1) ctas_per_rank = 32;
  dim3 grid(ctas_per_rank, emulate_all_ranks ? world : 1);
  const int br = emulate_all_ranks ? 0 : rank;
  if (is_bf16) p2p_allreduce_kernel<true><<<grid, P2P_THREADS, 0, stream>>>(c, br, (long)offset_bytes, (long)n, epoch, scale, two_shot);
  else p2p_allreduce_kernel<false><<<grid, P2P_THREADS, 0, stream>>>(c, br, (long)offset_bytes, (long)n, epoch, scale, two_shot);
  return rt1_check_launch("rt1_p2p_allreduce");
}
