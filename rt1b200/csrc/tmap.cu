// Host side of TMA: CUtensorMap encoding through the driver entry point (no -lcuda link dependency) with a small
// per-thread cache.  A training step issues ~300 TMA-fed launches whose operands sit at the same addresses step after
// step (the caching allocator hands the same blocks back), so each distinct (pointer, shape, box) is encoded once.
#include <unordered_map>

#include "sm100.cuh"

rt1_encode_tiled_fn rt1_get_encode_tiled() {
  static rt1_encode_tiled_fn fn = nullptr;
  if (fn) return fn;
  void* f = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) != cudaSuccess ||
      q != cudaDriverEntryPointSuccess)
    return nullptr;
  fn = (rt1_encode_tiled_fn)f;
  return fn;
}

namespace {
struct TmapKey {
  uint64_t v[8];
  bool operator==(const TmapKey& o) const {
    for (int i = 0; i < 8; ++i)
      if (v[i] != o.v[i]) return false;
    return true;
  }
};
struct TmapKeyHash {
  size_t operator()(const TmapKey& k) const {
    uint64_t h = 0x9E3779B97F4A7C15ull;
    for (int i = 0; i < 8; ++i) {
      h ^= k.v[i] + 0x9E3779B97F4A7C15ull + (h << 6) + (h >> 2);
      h *= 0xff51afd7ed558ccdull;
    }
    return (size_t)h;
  }
};
typedef std::unordered_map<TmapKey, CUtensorMap, TmapKeyHash> TmapCache;
// per thread: autograd runs the backward pass on its own thread, and a thread-local map needs no lock
TmapCache& tmap_cache() {
  static thread_local TmapCache c;
  return c;
}
constexpr size_t kTmapCacheMax = 8192;

int bind_context() {
  // The driver-API encoder needs a current context on the calling thread; autograd's backward threads may not
  // have touched the runtime yet, so bind the primary context once per thread.
  static thread_local bool ctx_bound = false;
  if (!ctx_bound) {
    RT1_CHECK_CUDA(cudaFree(0));
    ctx_bound = true;
  }
  return 0;
}
}  // namespace

int rt1_make_tmap_bf16_2d(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows,
                          uint32_t box_cols) {
  const TmapKey key = {{(uint64_t)(uintptr_t)base, rows, cols, ld, box_rows, box_cols, 2, 0}};
  TmapCache& cache = tmap_cache();
  auto it = cache.find(key);
  if (it != cache.end()) {
    *out = it->second;
    return 0;
  }
  if (int rc = bind_context()) return rc;
  rt1_encode_tiled_fn enc = rt1_get_encode_tiled();
  RT1_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled entry point unavailable");
  RT1_REQUIRE(((uintptr_t)base & 15) == 0 && (ld * 2) % 16 == 0, "TMA operand must be 16-byte aligned (base %p, ld %llu)",
              base, (unsigned long long)ld);
  cuuint64_t gdim[2] = {cols, rows};
  cuuint64_t gstr[1] = {ld * 2};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), gdim, gstr, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  RT1_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed with %d (rows %llu cols %llu ld %llu box %ux%u)", (int)r,
              (unsigned long long)rows, (unsigned long long)cols, (unsigned long long)ld, box_rows, box_cols);
  if (cache.size() >= kTmapCacheMax) cache.clear();
  cache.emplace(key, *out);
  return 0;
}

int rt1_make_tmap_bf16_nhwc(CUtensorMap* out, const void* base, uint64_t N, uint64_t H, uint64_t W, uint64_t C, uint32_t box_h,
                            uint32_t box_w, uint32_t box_c) {
  const TmapKey key = {{(uint64_t)(uintptr_t)base, N, H, W, C, ((uint64_t)box_h << 32) | box_w, box_c, 4}};
  TmapCache& cache = tmap_cache();
  auto it = cache.find(key);
  if (it != cache.end()) {
    *out = it->second;
    return 0;
  }
  if (int rc = bind_context()) return rc;
  rt1_encode_tiled_fn enc = rt1_get_encode_tiled();
  RT1_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled entry point unavailable");
  RT1_REQUIRE(((uintptr_t)base & 15) == 0 && C % 8 == 0 && box_c % 8 == 0 && box_h <= 256 && box_w <= 256 && box_c <= 256,
              "NHWC TMA operand: base %p C %llu box %ux%ux%u (16-byte alignment, box <= 256 per dimension)", base,
              (unsigned long long)C, box_h, box_w, box_c);
  cuuint64_t gdim[4] = {C, W, H, N};
  cuuint64_t gstr[3] = {C * 2, W * C * 2, H * W * C * 2};
  cuuint32_t box[4] = {box_c, box_w, box_h, 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), gdim, gstr, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  RT1_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled (NHWC) failed with %d (N %llu H %llu W %llu C %llu box %ux%ux%u)", (int)r,
              (unsigned long long)N, (unsigned long long)H, (unsigned long long)W, (unsigned long long)C, box_h, box_w, box_c);
  if (cache.size() >= kTmapCacheMax) cache.clear();
  cache.emplace(key, *out);
  return 0;
}
