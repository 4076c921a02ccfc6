// NVLink/NVSwitch peer-to-peer collectives fused with the optimizer (SURVEY §8 A1–A3).
//
// Replaces, on the one communication-bound hot path of RT-1 training, the whole chain
//     DDP bucket copy → ncclAllReduce → divide by world → torch.optim.Adam (foreach kernels) → zero_grad
// (reference distribute_train.py:99-110,231-238) with ONE kernel per gradient bucket that talks to the peers'
// memory directly.  Every rank's bf16 gradient and bf16 parameter buffers live in "symmetric" allocations that
// all peers have mapped; B200's NVSwitch gives each GPU full bandwidth to every peer, so a two-shot schedule is
// bandwidth-optimal without any ring:
//
//   barrier "ready"   : flag write to every peer (st.release.sys) + wait on own flags (ld.acquire.sys)
//   reduce-scatter    : rank r obtains the sum of shard r of the bucket
//                         P2P : PULLS the shard from every peer with 128-bit loads over NVLink, sums in fp32
//                         NVLS: one multimem.ld_reduce per 16 bytes — the switch adds the W copies in flight
//   optimizer         : x 1/world, bf16→fp32, AdamW on the fp32 master / moment shards that rank r owns
//   all-gather        : rank r writes the updated bf16 parameter shard into every rank's parameter buffer
//                         P2P : W stores to the mapped peer pointers;  NVLS: one multimem.st (switch multicast)
//   barrier "done"    : last CTA of the rank signals; afterwards each rank clears its own gradient bucket
//
// Bytes per direction per GPU for a bucket of N bytes: P2P 2N(W-1)/W, NVLS N(1+1/W) — NVLS wins from W = 4 up;
// each rank touches its fp32 state only for 1/W of the bucket (the sharded update SURVEY App. C argues for).
// The same file provides all-reduce-only kernels (one-shot for latency-bound sizes, two-shot / NVLS for
// bandwidth-bound ones) used for scalars and for the 1 KB–256 MB sweep.
//
// Latency hiding: a thread issues every load of its tile — peer (or multimem) gradient vectors AND the fp32
// master/moment vectors of the same elements — before it consumes any of them, one 512-thread CTA per SM: ≈17 MB
// outstanding per GPU against a bandwidth-delay product of ≈2.5 MB (NVLink) + ≈7 MB (HBM).  Measured on 2 x B200
// (profiles/r1_p2p_sweep_w2.json): 35.3 M parameters in 0.160 ms vs 0.340 ms for ncclAllReduce + fused Adam.
//
// Test/emulation mode: with gridDim.y == world the kernel plays ALL ranks in one launch on one GPU (rank =
// blockIdx.y, all CTAs co-resident), which is how the multi-rank logic is verified on a single B200 — separate
// launches that spin on each other on one GPU are not safe.
#include <string.h>

#include "adam.cuh"

#define P2P_MAX_WORLD 8
#define P2P_THREADS 512
#define P2P_INFLIGHT 16   // 16-byte peer loads a thread issues back to back

enum { P2P_ALGO_ONE_SHOT = 0, P2P_ALGO_TWO_SHOT = 1, P2P_ALGO_NVLS = 2 };

struct P2PComm {
  int world;
  void* grad[P2P_MAX_WORLD];        // bf16 gradient buffers (symmetric), indexed by rank
  void* param[P2P_MAX_WORLD];       // bf16 parameter buffers (symmetric)
  unsigned* flags[P2P_MAX_WORLD];   // signal pads: [2][P2P_MAX_WORLD] epochs (ready, done) + [1] CTA counter
  void* mc_grad;                    // multicast (NVLS) aliases of the two buffers, or null
  void* mc_param;
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
// Peer data is read once, after an acquire at system scope: L2-only loads (no L1 allocation, nothing stale).
__device__ __forceinline__ uint4 ld_peer_v4(const void* p) { return __ldcg(reinterpret_cast<const uint4*>(p)); }
__device__ __forceinline__ void st_peer_v4(void* p, uint4 v) { __stcg(reinterpret_cast<uint4*>(p), v); }
// NVLS: the switch sums the W copies of 8 bf16 (fp32 accumulation) / 4 fp32 and returns one vector
__device__ __forceinline__ uint4 mc_ld_reduce_bf16x8(const void* mc) {
  uint4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(mc) : "memory");
  return v;
}
__device__ __forceinline__ uint4 mc_ld_reduce_f32x4(const void* mc) {
  float4 f;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(f.x), "=f"(f.y), "=f"(f.z), "=f"(f.w) : "l"(mc) : "memory");
  return make_uint4(__float_as_uint(f.x), __float_as_uint(f.y), __float_as_uint(f.z), __float_as_uint(f.w));
}
__device__ __forceinline__ void mc_st_v4(void* mc, uint4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc), "f"(__uint_as_float(v.x)),
               "f"(__uint_as_float(v.y)), "f"(__uint_as_float(v.z)), "f"(__uint_as_float(v.w)) : "memory");
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

__device__ __forceinline__ void add_bf16x8(float* acc, const uint4& raw) {
  float f[8];
  unpack8(*reinterpret_cast<const bf16x8*>(&raw), f);
#pragma unroll
  for (int k = 0; k < 8; ++k) acc[k] += f[k];
}

// W = world size (compile time: the peer loop unrolls and the load batch is sized to it); NVLS selects multimem.
// A thread handles V output vectors (8 elements each) per tile and issues EVERY load of the tile — the SRC peer
// (or multimem) loads and the six fp32 state loads per vector — before it consumes any of them, so the NVLink round
// trip and the local HBM latency overlap instead of adding up (V x (SRC x 16 + 96) bytes in flight per thread).
//   prev_n > 0 : bucket [prev_off, prev_off + prev_n) was reduced by the previous launch of this stream; its "done"
//                barrier is folded into this launch's "ready" barrier (a rank that signals ready has, by stream
//                order, finished reading everybody's gradients of the previous bucket), after which the rank clears
//                its own copy of it — k buckets cost k + 1 barriers instead of 2k;
//   final      : run the closing barrier and clear this bucket too (last bucket of a step).
template <int W, bool NVLS>
__global__ void __launch_bounds__(P2P_THREADS, 1) p2p_allreduce_adamw_kernel(P2PComm c, P2PLocal loc, int base_rank, long off, long n,
                                                                            long prev_off, long prev_n, int final, unsigned epoch,
                                                                            AdamHyper h) {
  constexpr int SRC = NVLS ? 1 : W;                      // vectors fetched per output vector
  constexpr int V = (SRC * 4 + 24) * 2 <= 80 ? 2 : 1;    // output vectors per thread per tile (register budget)
  const int rank = base_rank + blockIdx.y;
  if (blockIdx.x == 0) signal_all(c, rank, 0, epoch);
  wait_all(c, rank, 0, epoch);
  if (prev_n > 0) {
    uint4* gz = reinterpret_cast<uint4*>(reinterpret_cast<bf16*>(c.grad[rank]) + prev_off);
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < (prev_n >> 3); i += (long)gridDim.x * blockDim.x)
      gz[i] = make_uint4(0, 0, 0, 0);
  }

  long lo, hi;
  shard_bounds(n, W, rank, lo, hi);
  const long base_el = off + lo;
  const float4* master = reinterpret_cast<const float4*>(loc.master[rank] + base_el);
  const float4* m = reinterpret_cast<const float4*>(loc.exp_avg[rank] + base_el);
  const float4* v = reinterpret_cast<const float4*>(loc.exp_avg_sq[rank] + base_el);
  const uint4* src[SRC];
  uint4* dst[SRC];
#pragma unroll
  for (int p = 0; p < SRC; ++p) {
    const int peer = (rank + p) % W;                     // start with the local copy, then round the peers
    src[p] = reinterpret_cast<const uint4*>(reinterpret_cast<const bf16*>(NVLS ? c.mc_grad : c.grad[peer]) + base_el);
    dst[p] = reinterpret_cast<uint4*>(reinterpret_cast<bf16*>(NVLS ? c.mc_param : c.param[peer]) + base_el);
  }
  const long nvec = (hi - lo) >> 3;
  // Outer batch: U output vectors per thread whose U x SRC gradient loads (NVLink or in-switch reduce: the long
  // latency) are ALL issued first; inner batches of V vectors then fetch their fp32 state from local HBM and update.
  // U = V: one inner batch per tile.  Deeper gradient-load batches (12-16 loads per thread issued up front) were measured
  // at 2 GPUs and were 25 % SLOWER for both schedules (register pressure: 128 registers with spills), so the outer batch
  // stays equal to the inner one; the loop structure is kept for tuning at other world sizes.
  constexpr int INFL = V * SRC;
  constexpr int U = (INFL / SRC) < V ? V : (INFL / SRC) / V * V;
  const long tile = (long)P2P_THREADS * U;
  for (long t0 = (long)blockIdx.x * tile; t0 < nvec; t0 += (long)gridDim.x * tile) {
    uint4 raw[U][SRC];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long i = t0 + u * P2P_THREADS + threadIdx.x;
      if (i < nvec) {
#pragma unroll
        for (int p = 0; p < SRC; ++p) raw[u][p] = NVLS ? mc_ld_reduce_bf16x8(src[p] + i) : ld_peer_v4(src[p] + i);
      }
    }
#pragma unroll
    for (int u0 = 0; u0 < U; u0 += V) {
      float4 st[V][6];
#pragma unroll
      for (int v_ = 0; v_ < V; ++v_) {
        const long i = t0 + (u0 + v_) * P2P_THREADS + threadIdx.x;
        if (i < nvec) {
          st[v_][0] = __ldcs(master + 2 * i); st[v_][1] = __ldcs(master + 2 * i + 1);
          st[v_][2] = __ldcs(m + 2 * i);      st[v_][3] = __ldcs(m + 2 * i + 1);
          st[v_][4] = __ldcs(v + 2 * i);      st[v_][5] = __ldcs(v + 2 * i + 1);
        }
      }
#pragma unroll
      for (int v_ = 0; v_ < V; ++v_) {
        const int u = u0 + v_;
        const long i = t0 + u * P2P_THREADS + threadIdx.x;
        if (i >= nvec) continue;
        float g[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
        for (int p = 0; p < SRC; ++p) add_bf16x8(g, raw[u][p]);
        float pp[8] = {st[v_][0].x, st[v_][0].y, st[v_][0].z, st[v_][0].w, st[v_][1].x, st[v_][1].y, st[v_][1].z, st[v_][1].w};
        float mm[8] = {st[v_][2].x, st[v_][2].y, st[v_][2].z, st[v_][2].w, st[v_][3].x, st[v_][3].y, st[v_][3].z, st[v_][3].w};
        float vv[8] = {st[v_][4].x, st[v_][4].y, st[v_][4].z, st[v_][4].w, st[v_][5].x, st[v_][5].y, st[v_][5].z, st[v_][5].w};
#pragma unroll
        for (int k = 0; k < 8; ++k) adam_update(pp[k], g[k], mm[k], vv[k], h);
        float4* mo = const_cast<float4*>(master) + 2 * i;
        float4* m1o = const_cast<float4*>(m) + 2 * i;
        float4* vo = const_cast<float4*>(v) + 2 * i;
        __stcs(mo, make_float4(pp[0], pp[1], pp[2], pp[3]));
        __stcs(mo + 1, make_float4(pp[4], pp[5], pp[6], pp[7]));
        __stcs(m1o, make_float4(mm[0], mm[1], mm[2], mm[3]));
        __stcs(m1o + 1, make_float4(mm[4], mm[5], mm[6], mm[7]));
        __stcs(vo, make_float4(vv[0], vv[1], vv[2], vv[3]));
        __stcs(vo + 1, make_float4(vv[4], vv[5], vv[6], vv[7]));
        const bf16x8 packed = pack8(pp);
        const uint4 pk = *reinterpret_cast<const uint4*>(&packed);
        if (NVLS) {
          mc_st_v4(dst[0] + i, pk);
        } else {
#pragma unroll
          for (int p = 0; p < SRC; ++p) st_peer_v4(dst[p] + i, pk);
        }
      }
    }
  }
  if (!final) {
    __threadfence_system();   // parameter pushes are performed system-wide before this launch can be seen as finished
    return;
  }
  rank_done_signal(c, rank, epoch);
  wait_all(c, rank, 1, epoch);
  // every peer has consumed this rank's gradients: clear the bucket for the next step
  uint4* gz = reinterpret_cast<uint4*>(reinterpret_cast<bf16*>(c.grad[rank]) + off);
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < (n >> 3); i += (long)gridDim.x * blockDim.x)
    gz[i] = make_uint4(0, 0, 0, 0);
}

// In-place sum all-reduce of a symmetric buffer (bf16 or fp32), scaled by `scale`.
//   TWO_SHOT : reduce-scatter by pulling + all-gather by pushing (bandwidth-optimal P2P schedule on NVSwitch)
//   NVLS     : multimem.ld_reduce of the rank's shard + multimem.st of the result (in-switch reduce + multicast)
//   ONE_SHOT : every rank pulls everything and writes only its own copy (one barrier less per byte; wins below a
//              few hundred KB where the barriers dominate).  Every rank reads all of [0, n) from every peer, so
//              nobody may overwrite the inputs while a peer could still be reading them — the result goes to
//              elements [n, 2n) of the rank's own buffer instead.
template <int W, bool IS_BF16, int ALGO>
__global__ void __launch_bounds__(P2P_THREADS, 1) p2p_allreduce_kernel(P2PComm c, int base_rank, long off_bytes, long n, unsigned epoch,
                                                                      float scale) {
  constexpr int EPV = IS_BF16 ? 8 : 4;  // elements per 16-byte vector
  constexpr bool NVLS = ALGO == P2P_ALGO_NVLS;
  constexpr int SRC = NVLS ? 1 : W;
  constexpr int U = P2P_INFLIGHT / SRC < 1 ? 1 : P2P_INFLIGHT / SRC;
  const int rank = base_rank + blockIdx.y;
  if (blockIdx.x == 0) signal_all(c, rank, 0, epoch);
  wait_all(c, rank, 0, epoch);
  long lo = 0, hi = n;
  if (ALGO != P2P_ALGO_ONE_SHOT) {
    long chunk = ((n + W - 1) / W + EPV - 1) / EPV * EPV;
    lo = min(n, chunk * rank);
    hi = min(n, lo + chunk);
  }
  const size_t esz = IS_BF16 ? 2 : 4;
  const uint4* src[SRC];
  uint4* dst[SRC];
#pragma unroll
  for (int p = 0; p < SRC; ++p) {
    const int peer = (rank + p) % W;
    char* b = reinterpret_cast<char*>(NVLS ? c.mc_grad : c.grad[peer]) + off_bytes + lo * esz;
    src[p] = reinterpret_cast<const uint4*>(b);
    dst[p] = reinterpret_cast<uint4*>(b);
  }
  uint4* own_out = reinterpret_cast<uint4*>(reinterpret_cast<char*>(c.grad[rank]) + off_bytes + n * esz);   // one-shot result
  const long nvec = (hi - lo) / EPV;
  const long tile = (long)P2P_THREADS * U;
  for (long t0 = (long)blockIdx.x * tile; t0 < nvec; t0 += (long)gridDim.x * tile) {
    uint4 raw[U][SRC];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long i = t0 + u * P2P_THREADS + threadIdx.x;
      if (i < nvec) {
#pragma unroll
        for (int p = 0; p < SRC; ++p)
          raw[u][p] = NVLS ? (IS_BF16 ? mc_ld_reduce_bf16x8(src[p] + i) : mc_ld_reduce_f32x4(src[p] + i)) : ld_peer_v4(src[p] + i);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long i = t0 + u * P2P_THREADS + threadIdx.x;
      if (i >= nvec) continue;
      float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
      for (int p = 0; p < SRC; ++p) {
        if (IS_BF16) {
          add_bf16x8(acc, raw[u][p]);
        } else {
          acc[0] += __uint_as_float(raw[u][p].x); acc[1] += __uint_as_float(raw[u][p].y);
          acc[2] += __uint_as_float(raw[u][p].z); acc[3] += __uint_as_float(raw[u][p].w);
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
      if (ALGO == P2P_ALGO_ONE_SHOT) {
        st_peer_v4(own_out + i, outv);
      } else if (NVLS) {
        mc_st_v4(dst[0] + i, outv);
      } else {
#pragma unroll
        for (int p = 0; p < SRC; ++p) st_peer_v4(dst[p] + i, outv);
      }
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

static int p2p_fill(P2PComm& c, int world, void* const* grad, void* const* param, void* const* flags, void* mc_grad, void* mc_param) {
  RT1_REQUIRE(world >= 1 && world <= P2P_MAX_WORLD, "p2p: world size %d not in [1, %d]", world, P2P_MAX_WORLD);
  c.world = world;
  for (int i = 0; i < P2P_MAX_WORLD; ++i) {
    c.grad[i] = i < world ? grad[i] : nullptr;
    c.param[i] = (i < world && param) ? param[i] : nullptr;
    c.flags[i] = i < world ? (unsigned*)flags[i] : nullptr;
  }
  c.mc_grad = mc_grad;
  c.mc_param = mc_param;
  return 0;
}

#define P2P_FOR_WORLD(W_, ...) \
  switch (W_) {                \
    case 1: { constexpr int W = 1; __VA_ARGS__; } break; \
    case 2: { constexpr int W = 2; __VA_ARGS__; } break; \
    case 3: { constexpr int W = 3; __VA_ARGS__; } break; \
    case 4: { constexpr int W = 4; __VA_ARGS__; } break; \
    case 5: { constexpr int W = 5; __VA_ARGS__; } break; \
    case 6: { constexpr int W = 6; __VA_ARGS__; } break; \
    case 7: { constexpr int W = 7; __VA_ARGS__; } break; \
    default: { constexpr int W = 8; __VA_ARGS__; } break; \
  }

// grad/param/flags: host arrays of `world` device pointers (peer-mapped).  master/exp_avg/exp_avg_sq: host arrays
// of device pointers; outside emulation only entry [rank] is used.  mc_grad/mc_param: multicast aliases of the
// gradient/parameter buffers (both non-null selects the NVLS schedule), else null.  [prev_offset, prev_offset+prev_n)
// is the bucket reduced by the PREVIOUS call of the step (prev_n = 0: none) — it is cleared here, after this call's
// opening barrier; final = 1 marks the last bucket of a step (closing barrier, this bucket cleared too).  A lone
// call passes prev_n = 0, final = 1.  emulate_all_ranks = 1 plays
// every rank in one launch on the current device (single-GPU test mode; needs world * ctas_per_rank co-resident
// CTAs; P2P schedule only).
RT1_API int rt1_p2p_allreduce_adamw(int world, int rank, void* const* grad, void* const* param, void* const* flags,
                                    float* const* master, float* const* exp_avg, float* const* exp_avg_sq, void* mc_grad,
                                    void* mc_param, long long offset, long long n, long long prev_offset, long long prev_n,
                                    int final, unsigned epoch, float lr, float beta1, float beta2, float eps,
                                    float weight_decay, float bias_corr1, float bias_corr2, int ctas_per_rank,
                                    int emulate_all_ranks, cudaStream_t stream) {
  P2PComm c;
  int rc = p2p_fill(c, world, grad, param, flags, mc_grad, mc_param);
  if (rc) return rc;
  RT1_REQUIRE(n % 8 == 0 && offset % 8 == 0 && prev_n % 8 == 0 && prev_offset % 8 == 0,
              "p2p_allreduce_adamw: offset/length must be multiples of 8 elements");
  const bool nvls = mc_grad != nullptr && mc_param != nullptr;
  RT1_REQUIRE(!(nvls && emulate_all_ranks), "p2p_allreduce_adamw: the NVLS schedule cannot be emulated on one GPU");
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
  const int br = emulate_all_ranks ? 0 : rank;
  P2P_FOR_WORLD(world,
                if (nvls) p2p_allreduce_adamw_kernel<W, true><<<grid, P2P_THREADS, 0, stream>>>(c, loc, br, (long)offset, (long)n, (long)prev_offset, (long)prev_n, final, epoch, h);
                else p2p_allreduce_adamw_kernel<W, false><<<grid, P2P_THREADS, 0, stream>>>(c, loc, br, (long)offset, (long)n, (long)prev_offset, (long)prev_n, final, epoch, h));
  return rt1_check_launch("rt1_p2p_allreduce_adamw");
}

template <int W, bool IS_BF16>
static void launch_allreduce(int algo, dim3 grid, cudaStream_t stream, const P2PComm& c, int br, long off, long n, unsigned epoch, float scale) {
  if (algo == P2P_ALGO_NVLS) p2p_allreduce_kernel<W, IS_BF16, P2P_ALGO_NVLS><<<grid, P2P_THREADS, 0, stream>>>(c, br, off, n, epoch, scale);
  else if (algo == P2P_ALGO_TWO_SHOT) p2p_allreduce_kernel<W, IS_BF16, P2P_ALGO_TWO_SHOT><<<grid, P2P_THREADS, 0, stream>>>(c, br, off, n, epoch, scale);
  else p2p_allreduce_kernel<W, IS_BF16, P2P_ALGO_ONE_SHOT><<<grid, P2P_THREADS, 0, stream>>>(c, br, off, n, epoch, scale);
}

// In-place all-reduce (sum * scale) of `n` elements at byte offset `offset_bytes` of the symmetric buffers `buf`.
// algo: 0 one-shot (result goes to elements [n, 2n) of the rank's own buffer), 1 two-shot P2P, 2 NVLS (needs `mc_buf`,
// the multicast alias of `buf`).
RT1_API int rt1_p2p_allreduce(int world, int rank, void* const* buf, void* const* flags, void* mc_buf, long long offset_bytes,
                              long long n, int is_bf16, float scale, int algo, unsigned epoch, int ctas_per_rank,
                              int emulate_all_ranks, cudaStream_t stream) {
  P2PComm c;
  int rc = p2p_fill(c, world, buf, nullptr, flags, mc_buf, nullptr);
  if (rc) return rc;
  RT1_REQUIRE(n % (is_bf16 ? 8 : 4) == 0 && offset_bytes % 16 == 0, "p2p_allreduce: 16-byte granularity required");
  RT1_REQUIRE(algo >= 0 && algo <= 2, "p2p_allreduce: unknown algorithm %d", algo);
  RT1_REQUIRE(algo != P2P_ALGO_NVLS || (mc_buf != nullptr && !emulate_all_ranks), "p2p_allreduce: NVLS needs a multicast mapping and real ranks");
  if (ctas_per_rank < 1) ctas_per_rank = 32;
  dim3 grid(ctas_per_rank, emulate_all_ranks ? world : 1);
  const int br = emulate_all_ranks ? 0 : rank;
  P2P_FOR_WORLD(world,
                if (is_bf16) launch_allreduce<W, true>(algo, grid, stream, c, br, (long)offset_bytes, (long)n, epoch, scale);
                else launch_allreduce<W, false>(algo, grid, stream, c, br, (long)offset_bytes, (long)n, epoch, scale));
  return rt1_check_launch("rt1_p2p_allreduce");
}
