/* rt1b200 — C ABI of the sm_100a kernel library (rt1b200/lib/librt1b200.so).
 *
 * Conventions: every function returns 0 on success, otherwise a CUDA error code or -1 (argument/shape error) and
 * rt1_last_error() holds a message (thread-local).  All pointers are device pointers unless stated otherwise;
 * bf16 tensors are passed as void*; every launch goes to the given stream and returns without synchronising.
 * No compute entry point allocates device memory: scratch buffers are caller-provided (their sizes are given per call).
 * The one allocating call is rt1_symm_alloc (cudaMalloc of a peer-mappable buffer at set-up time).
 *
 * "Reference" citations name the file:line of tanhuajie/Pytorch-RT1-for-Distributed-Training whose behaviour
 * the kernel reproduces (the reference itself has no native code; it reaches these ops through ATen/cuDNN/
 * cuBLAS/NCCL).
 */
#ifndef RT1B200_H_
#define RT1B200_H_

#include <cuda_runtime.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- library ------------------------------------------------------------------------------------------- */
const char* rt1_last_error(void);
int rt1_abi_version(void);
int rt1_device_info(int* sm_count, int* cc_major, int* cc_minor, long long* smem_optin);
/* Register a caller-owned uint32 in device memory whose value is added to every dropout seed on the current device (NULL
 * clears it).  A captured CUDA graph bakes the seeds passed by value; advancing this word from a node of the graph gives
 * every replay fresh dropout masks (forward and backward of one replay read the same value). */
int rt1_set_seed_epoch(const void* dev_u32);
/* Publish a device float to pinned host memory from inside a captured graph: *dst_host = *src_dev, then ++*counter_dev is
 * written to *flag_host (both host pointers must be pinned, device-accessible memory).  The host polls *flag_host. */
int rt1_publish_scalar(const float* src_dev, float* dst_host, unsigned* counter_dev, unsigned* flag_host, cudaStream_t stream);
/* Benchmark helper: stream `bytes` of writes through `buf` (use a buffer larger than the 126 MB L2). */
int rt1_l2_flush(void* buf, long long bytes, unsigned tag, cudaStream_t stream);

/* ---- GEMM: tcgen05.mma + TMEM accumulator + TMA operand loads ---------------------------------------------
 * out[M,N] = alpha * op(A) . op(B)^T (+ bias[N]) -> dropout(p, seed) -> (+ residual[M,N]).
 * a_mn = 0: A is [M,K] row-major (lda);  a_mn = 1: A is stored [K,M] row-major (a transposed view, free).
 * b_mn = 0: B is [N,K] row-major (ldb);  b_mn = 1: B is stored [K,N] row-major.
 * out is bf16 (out_fp32 = 0) or fp32; split_k > 1 / accumulate need fp32 `out` (atomic accumulation; zero it first).
 * col_sum / col_sumsq (optional, fp32 [N], caller-zeroed): per-column sum and sum of squares of the fp32 result
 * (BatchNorm batch statistics fused into the producing GEMM).
 * split_k > 1 or accumulate exclude bias, residual, dropout and column statistics (rejected with an error).
 * Replaces: nn.Linear of the decoder (reference transformer.py:45-49,126,171-173) and every 1x1 convolution of
 * the image tokenizer (film_efficientnet_encoder.py:185-224,329-336; pretrained_efficientnet_encoder.py:45-51). */
int rt1_gemm_bf16(const void* A, long long lda, int a_mn, const void* B, long long ldb, int b_mn, void* out,
                  long long ldo, int out_fp32, int M, int N, int K, const void* bias, const void* residual,
                  long long ldr, float alpha, float dropout_p, unsigned seed, int split_k, int accumulate,
                  float* col_sum, float* col_sumsq, cudaStream_t stream);
/* BatchNorm "tail" of a statistics-producing kernel: the CTA that adds the LAST contribution to sum / sumsq (a ticket counter,
 * no CTA waits for another) turns them into a = gamma*rstd, b = beta - mean*a, mean, rstd and updates the running statistics —
 * what rt1_bn_finalize does as a launch of its own.  ticket: one uint32, zeroed by the caller before every use. */
typedef struct Rt1BnTail {
  const void* gamma;        /* bf16 [C] */
  const void* beta;         /* bf16 [C] */
  float* running_mean;      /* fp32 [C] or NULL */
  float* running_var;
  float* a;                 /* fp32 [C] outputs */
  float* b;
  float* mean;
  float* rstd;
  unsigned int* ticket;
  float count;              /* elements per channel */
  float momentum;
  float eps;
} Rt1BnTail;
/* rt1_gemm_bf16 with bf16 output, no epilogue terms, column statistics AND their BatchNorm tail (host pointer to the struct) */
int rt1_gemm_bf16_bn(const void* A, long long lda, int a_mn, const void* B, long long ldb, int b_mn, void* out, long long ldo,
                     int M, int N, int K, float* col_sum, float* col_sumsq, const Rt1BnTail* tail, cudaStream_t stream);
int rt1_colsum_bf16(const void* x, long long ld, int M, int N, void* out_bf16, cudaStream_t stream);
int rt1_cast_f32_bf16(const float* in, void* out_bf16, long long n, cudaStream_t stream);
int rt1_dropout_apply(const void* in, void* out, long long n, float p, unsigned seed, cudaStream_t stream);

/* ---- LayerNorm (reference transformer.py:123,129,132-141) ------------------------------------------------ */
int rt1_layernorm_fwd(const void* x, const void* w, const void* b, void* y, float* mean, float* rstd, int rows,
                      int D, float eps, cudaStream_t stream);
long long rt1_layernorm_bwd_scratch_floats(int rows, int D);
/* scratch: fp32 [rt1_layernorm_bwd_scratch_floats(rows, D)] of per-CTA dgamma/dbeta partials */
int rt1_layernorm_bwd(const void* x, const void* w, const void* dy, const float* mean, const float* rstd, void* dx,
                      void* dw, void* db, float* scratch, int rows, int D, cudaStream_t stream);
/* Same computation for hosts that cannot provide `scratch` (all device memory reserved at start-up or owned by another
 * allocator, so no cudaMalloc may happen while the robot runs): the partial sums go to a fixed __device__ array inside
 * the library (1.2 MB per GPU for the life of the process).  This entry point has no counterpart in the reference.
 * Limits: D in {256, 512, 1024} (rt1_layernorm_bwd_noalloc_max_d() == 1024; wider rows are rejected with an error),
 * any number of rows.  It performs no allocation.  The first call on each device looks up the array's device address
 * (cudaGetSymbolAddress, cached per device): make that first call before starting any stream capture.  All calls on a
 * device share the array: calls whose kernels overlap on the GPU race on it — keep them on one stream or order them
 * with events.  Nothing in the array needs clearing between calls. */
int rt1_layernorm_bwd_noalloc(const void* x, const void* w, const void* dy, const float* mean, const float* rstd,
                              void* dx, void* dw, void* db, int rows, int D, cudaStream_t stream);
long long rt1_layernorm_bwd_noalloc_max_d(void);

/* ---- fused masked attention over the RT-1 token window (reference transformer.py:82-109) ------------------
 * qkv [B,S,3,H,DK] bf16 -> out [B,S,H*DK]; mask uint8 [S,S] (0 = blocked) or NULL.
 * S <= 128 with DK in {64, 128} (RT-1: S = 66, DK = 128) runs on tcgen05 tensor cores: one CTA per (batch, head), TMA-loaded
 * Q/K/V/dO tiles, scores and all five backward contractions accumulated in TMEM, softmax/mask/dropout thread-local on TMEM
 * lanes (csrc/attention_tc.cu).  Other shapes (DK % 8 == 0, DK <= 256, S <= 160) use the generic CUDA-core kernels. */
int rt1_attention_fwd(const void* qkv, const unsigned char* mask, void* out, int B, int S, int H, int DK, float p_drop,
                      unsigned seed, cudaStream_t stream);
int rt1_attention_bwd(const void* qkv, const unsigned char* mask, const void* dout, void* dqkv, int B, int S, int H,
                      int DK, float p_drop, unsigned seed, cudaStream_t stream);

/* ---- action-token cross-entropy (reference transformer_network.py:304-322) -------------------------------
 * logits [R,A,V] bf16, labels int64 [R,A] -> loss[r] = mean_a CE / num_items, pred = argmax.  row_ce / row_lse: fp32 [R*A]. */
int rt1_action_ce_fwd(const void* logits, const long long* labels, float* loss, float* row_ce, float* row_lse,
                      long long* pred, int R, int A, int V, float num_items, cudaStream_t stream);
int rt1_action_ce_bwd(const void* logits, const long long* labels, const float* row_lse, const float* dloss,
                      void* dlogits, int R, int A, int V, float num_items, cudaStream_t stream);

/* ---- frame pre-processing (reference film_efficientnet/preprocessors.py:37-56) ---------------------------
 * in [N,3,H,W] uint8|f32 -> out [N,H,W,8] bf16 NHWC with the pad-and-shift crop (one offset per batch). */
int rt1_preprocess_nhwc(const void* in, int in_is_uint8, void* out, int N, int H, int W, int shift_h, int shift_w,
                        int pad_h, int pad_w, cudaStream_t stream);
/* same, the (dh, dw) offset read from device memory (int32[2]) at execution time — for CUDA-graph-captured steps */
int rt1_preprocess_nhwc_dev(const void* in, int in_is_uint8, void* out, int N, int H, int W, const int* shift_dev, int pad_h,
                            int pad_w, cudaStream_t stream);

/* ---- optimizer: flat fused AdamW (reference distribute_train.py:100 torch.optim.Adam) ---------------------- */
int rt1_fused_adamw(void* param_bf16, float* master, void* grad, int grad_is_bf16, float* exp_avg, float* exp_avg_sq,
                    long long n, float lr, float beta1, float beta2, float eps, float weight_decay, float bias_corr1,
                    float bias_corr2, float grad_scale, int zero_grad, cudaStream_t stream);

/* ---- TokenLearner (reference tokenizers/token_learner.py:64-95); C % 64 == 0, T <= 8, P <= 512 ----------------------
 * hpre [n*P,64] = LayerNorm(feats) . W1^T + b1 comes from rt1_layernorm_fwd + rt1_gemm_bf16; these finish the module:
 * softmax over the P positions of gelu(hpre).W2^T + b2 and the weighted pooling (fwd); dwmap → softmax backward → dhpre
 * plus dW2/db2 sums (bwd; scratch fp32 [T*64+T], zeroed by the call); dfeats += wmap.dtok (direct pooling gradient). */
int rt1_tl_pool_fwd(const void* feats, const void* hpre, const void* w2, const void* b2, void* tokens, long long out_stride,
                    float* wmap, int n, int P, int C, int T, cudaStream_t stream);
int rt1_tl_pool_bwd(const void* feats, const void* hpre, const float* wmap, const void* w2, const void* dtokens,
                    long long dtok_stride, void* dhpre, float* scratch, int n, int P, int C, int T, cudaStream_t stream);
int rt1_tl_direct_bwd(const float* wmap, const void* dtokens, long long dtok_stride, void* dfeats, int n, int P, int C, int T,
                      cudaStream_t stream);

/* ---- image tokenizer, NHWC bf16 (reference film_efficientnet_encoder.py:142-244,272-336; film_conditioning_layer.py) ----
 * Activations are [N, H*W, C] bf16 with C % 8 == 0; BatchNorm runs on batch statistics (training) that the PRODUCING kernel
 * emits (sum / sumsq, fp32 [C], caller-zeroed): a = gamma*rstd, b = beta - mean*a are what consumers apply on load. */
int rt1_bn_finalize(const float* sum, const float* sumsq, float count, const void* gamma, const void* beta, float* running_mean,
                    float* running_var, float momentum, float eps, int training, float* a, float* b, float* mean, float* rstd,
                    int C, cudaStream_t stream);
/* pooled[n,c] += sum_hw SiLU(a x + b)   (squeeze of squeeze-excite; caller-zeroed) */
int rt1_bn_act_pool(const void* x, const float* a, const float* b, float* pooled, int N, int HW, int C, cudaStream_t stream);
/* y = act(a x + b) * s[n,c]   (s = SE scale or NULL; silu = 0 leaves the affine only) */
int rt1_bn_act_scale(const void* x, const float* a, const float* b, const float* s, void* y, int N, int HW, int C, int silu,
                     cudaStream_t stream);
/* out = FiLM([skip +] sd[n] * (a y + b)): stochastic depth ("row" mode) scales the BRANCH, the skip is added unscaled;
 * gamma/beta rows are taken from gb[n, gb_off .. gb_off+2C) (no FiLM when gb = NULL) */
int rt1_block_finalize(const void* y, const float* a, const float* b, const void* skip, const float* sd, const void* gb,
                       long long gb_stride, long long gb_off, void* out, int N, int HW, int C, cudaStream_t stream);
int rt1_block_finalize_bwd(const void* dout, const void* y, const void* skip, const void* gb, long long gb_stride, long long gb_off,
                           const float* sd, const float* a, const float* mean, const float* rstd, const void* gamma3,
                           const void* beta3, float* S, float* c1, float* c2, void* dgb, void* dgamma3, void* dbeta3, void* dy,
                           void* dskip, int N, int HW, int C, cudaStream_t stream);
/* ds[n,c] = sum_hw g * SiLU(a x + b)   (gradient of the SE scale) */
int rt1_se_scale_bwd_reduce(const void* g, const void* x, const float* a, const float* b, float* ds, int N, int HW, int C,
                            cudaStream_t stream);
/* ONE pass over (g, x) for both the SE-scale gradient and the separable parts of the BatchNorm-backward sums of the same
 * tensor: out5 fp32 [5][N][C] = ds, A1..A4 (zeroed by the call); rt1_se_bwd combines A1..A4 with s and dpool (known only
 * after the SE backward) into the [2,C] sums that rt1_bn_act_bwd accepts with sums_ready = 1. */
int rt1_se_bn_bwd_reduce(const void* g, const void* x, const float* a, const float* b, const float* mean, const float* rstd,
                         float* out5, int N, int HW, int C, cudaStream_t stream);
/* backward of t = act(BN(x)) [* s + pooled gradient]: reduce pass into sums fp32 [2,C] (skipped when sums_ready: the depthwise
 * data-gradient kernel or rt1_se_bwd filled them), then one apply pass that writes dx and dgamma / dbeta (16-byte aligned) */
int rt1_bn_act_bwd(const void* g, const void* x, const float* a, const float* b, const float* mean, const float* rstd,
                   const float* s, const float* dpool, float* sums, int sums_ready, void* dgamma, void* dbeta, void* dx, int N,
                   int HW, int C, int silu, cudaStream_t stream);
/* squeeze-excite FC pair on per-frame vectors (z1, s fp32 outputs; Cs even, <= 128) and its backward + weight gradients */
int rt1_se_fc_fwd(const float* pooled_sum, float inv_hw, const void* w1, const void* b1, const void* w2, const void* b2, float* z1,
                  float* s, int N, int C, int Cs, cudaStream_t stream);
/* Backward of the FC pair, its parameter gradients and (parts != NULL) the BatchNorm-backward sums of the scaled tensor, as two
 * launches in which a CTA owns a 32-channel slice for all frames.  dhid: fp32 [N, ceil4(Cs)] scratch ZEROED BY THE CALLER;
 * dpooled fp32 [N,C] is the per-pixel gradient (already / HW); parts = out5 of rt1_se_bn_bwd_reduce, sums fp32 [2,C] out. */
int rt1_se_bwd(const float* ds, const float* s, const float* z1, const float* pooled_sum, float inv_hw, const void* w1, const void* w2,
               const float* parts, float* dhid, float* dpooled, float* sums, void* dw1, void* db1, void* dw2, void* db2, int N, int C,
               int Cs, cudaStream_t stream);
/* depthwise k x k (k in {3,5}, stride in {1,2}, pad (k-1)/2): forward (+ BN statistics of the output), data gradient
 * (x SiLU'(BN(x_raw)) + BN-backward sums when x_raw/a_in/... are given), weight gradient (scratch fp32 [k*k*C]) */
int rt1_dwconv_fwd(const void* x, const float* a_in, const float* b_in, const void* w, void* y, float* sum, float* sumsq, int N,
                   int H, int W, int C, int k, int stride, cudaStream_t stream);
/* rt1_dwconv_fwd + the BatchNorm tail of its output statistics */
int rt1_dwconv_fwd_bn(const void* x, const float* a_in, const float* b_in, const void* w, void* y, float* sum, float* sumsq,
                      const Rt1BnTail* tail, int N, int H, int W, int C, int k, int stride, cudaStream_t stream);
int rt1_dwconv_bwd_data(const void* dout, const void* w, const void* x_raw, const float* a_in, const float* b_in, const float* mean,
                        const float* rstd, void* din, float* sums, int N, int H, int W, int C, int k, int stride,
                        cudaStream_t stream);
int rt1_dwconv_wgrad(const void* dout, const void* x, const float* a_in, const float* b_in, float* scratch, void* dw, int N, int H,
                     int W, int C, int k, int stride, cudaStream_t stream);
/* stem 3x3 stride 2 on the [N,H,W,8] frames produced by rt1_preprocess_nhwc (channels 0..2 used): */
/* im2col of the stem window: cols bf16 [N*Ho*Wo, 32], column ci*9+ky*3+kx (27 used) — the stem then runs as rt1_gemm_bf16 */
int rt1_stem_im2col(const void* x, void* cols, int N, int H, int W, cudaStream_t stream);

/* ---- action (de)tokenisation (reference tokenizers/action_tokenizer.py:105-159) -------------------------------------------
 * vals fp32 [R, A] (one column per token; discrete values as floats) <-> tokens int64 [R, A]; low/high fp32 [A] (for a
 * discrete token high = n), discrete uint8 [A].  Box: clamp, affine to [0, vocab-1], int32 truncation; Discrete: identity,
 * and on the way back `token > n ? 0 : token`. */
int rt1_action_tokenize(const float* vals, const float* low, const float* high, const unsigned char* discrete, long long* tokens,
                        long long R, int A, int vocab, cudaStream_t stream);
int rt1_action_detokenize(const long long* tokens, const float* low, const float* high, const unsigned char* discrete, float* vals,
                          long long R, int A, int vocab, cudaStream_t stream);

/* ---- synthetic Language-Table frames (replaces the TinyRenderer call of reference language_table.py:579-597 and the
 *      crop + resize of eval/wrappers.py:88-96; oracle = the numpy restatement in rt1b200/envs/raster.py) ----------------
 * tris fp32 [E, max_tris, 12] screen-space triangles (3 x (x,y,z) + rgb, stride 12 floats), ntris int32 [E] → uint8 [E,H,W,3] */
int rt1_raster(const float* tris, const int* ntris, int max_tris, unsigned char* out, int E, int H, int W, float bg_r, float bg_g,
               float bg_b, cudaStream_t stream);
int rt1_crop_resize(const unsigned char* in, float* out, int E, int H, int W, int OH, int OW, float crop, cudaStream_t stream);
/* training input pipeline: per-frame crop box (top, left, height, width in source pixels; fp32 [E,4]) + bilinear resize of raw
 * uint8 HWC frames to [E,3,OH,OW] (uint8, or fp32 in [0,1]) — replaces the per-frame PIL crop+resize of the reference's data
 * loader workers (load_np_dataset.py:8-39) */
int rt1_crop_resize_boxes(const unsigned char* in, const float* boxes, void* out, int out_uint8, int E, int H, int W, int OH, int OW,
                          cudaStream_t stream);

/* ---- NVLink / NVSwitch collectives fused with the optimizer (replaces DDP + NCCL all-reduce + Adam,
 *      reference distribute_train.py:99-110,231-238) ------------------------------------------------------------
 * Symmetric buffers: every rank allocates the same sizes in the same order and maps every peer's allocation
 * (rt1_symm_alloc / rt1_symm_open exchange cudaIpc handles; any allocator that yields peer-mapped pointers — and,
 * for NVLS, a multicast alias — can be used instead).  `grad`, `param`, `flags`, `buf` are HOST arrays of `world`
 * device pointers indexed by rank; `flags` entries are zero-initialised pads of rt1_p2p_flag_words() uint32.
 * `epoch` must increase by one per collective, identically on all ranks. */
int rt1_symm_alloc(long long bytes, void** ptr, unsigned char* handle64);
int rt1_symm_open(const unsigned char* handle64, void** ptr);
int rt1_symm_close(void* ptr);
int rt1_symm_free(void* ptr);
int rt1_p2p_flag_words(void);
/* One kernel per bucket [offset, offset+n) of the flat buffers: ready barrier, reduce-scatter (P2P pulls, or
 * multimem.ld_reduce when mc_grad/mc_param are non-null), 1/world scale, AdamW on the rank's fp32 shard, all-gather of
 * the bf16 parameters (P2P pushes or multimem.st), done barrier, gradient bucket cleared.  master/exp_avg/exp_avg_sq
 * are host arrays of device pointers of which only [rank] is read (all of them with emulate_all_ranks = 1, the
 * single-GPU test mode that plays every rank in one launch).  Buckets of one step chain: [prev_offset, +prev_n) is the
 * bucket of the previous call (0 = none), cleared here after this call's opening barrier; final = 1 on the last bucket
 * of the step runs the closing barrier and clears this bucket too.  A lone call passes prev_n = 0, final = 1. */
int rt1_p2p_allreduce_adamw(int world, int rank, void* const* grad, void* const* param, void* const* flags,
                            float* const* master, float* const* exp_avg, float* const* exp_avg_sq, void* mc_grad,
                            void* mc_param, long long offset, long long n, long long prev_offset, long long prev_n,
                            int final, unsigned epoch, float lr, float beta1, float beta2, float eps, float weight_decay,
                            float bias_corr1, float bias_corr2, int ctas_per_rank, int emulate_all_ranks,
                            cudaStream_t stream);
/* In-place sum * scale of n elements (bf16 or fp32).  algo 0 = one-shot (result lands in elements [n, 2n) of the
 * rank's own buffer), 1 = two-shot P2P, 2 = NVLS (needs mc_buf, the multicast alias of buf). */
int rt1_p2p_allreduce(int world, int rank, void* const* buf, void* const* flags, void* mc_buf, long long offset_bytes,
                      long long n, int is_bf16, float scale, int algo, unsigned epoch, int ctas_per_rank,
                      int emulate_all_ranks, cudaStream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* RT1B200_H_ */
