/* C ABI of libkubedl_b200.so — the hand-written sm_100a kernel library of the kubedl_b200
 * worker runtime.  Every entry point returns 0 (KDL_OK) on success; on failure it returns a
 * non-zero code and kdl_last_error() describes what went wrong.  Pointers are device pointers
 * unless the name says otherwise; `stream` is a cudaStream_t passed as void*-compatible handle.
 *
 * The reference (jiaqianjing/kubedl) is a pure-Go operator with no device code at all
 * (SURVEY.md §2.2); each block below cites the stock-worker behaviour it replaces and the
 * reference file that wires that worker up.
 */
#ifndef KUBEDL_B200_H_
#define KUBEDL_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __CUDACC__
#include <cuda_runtime.h>
#else
typedef struct CUstream_st* cudaStream_t;
#endif

#ifdef __cplusplus
extern "C" {
#endif

#define KDL_MAX_WORLD 8
#define KDL_AR_MAX_BLOCKS 512

/* dtype codes */
#define KDL_BF16 0
#define KDL_F32 1

const char* kdl_last_error(void);
int kdl_version(void);

/* ---- N8: fused BatchNorm(+residual)(+ReLU), NHWC bf16 viewed as [M, C] ------------------
 * Replaces cuDNN BN + ReLU + add of the stock PyTorch worker image
 * (reference example/pytorch/pytorch_job_mnist_mpi.yaml:15).  `workspace` is caller-owned
 * scratch of kdl_bn_workspace_bytes(M, C) bytes whose first 256 bytes must be zero before the
 * first use (the kernels restore them to zero). */
size_t kdl_bn_workspace_bytes(int64_t M, int C);
int kdl_bn_fwd(const void* x, const void* res, void* y, int64_t M, int C, const float* gamma,
               const float* beta, float* running_mean, float* running_var, float eps,
               float momentum, int relu, float* save_mean_rstd, void* workspace,
               size_t workspace_bytes, cudaStream_t stream);
/* relu != 0: the ReLU mask comes from y, or — y == NULL and beta given, layers without a residual input —
 * is recomputed from x with the forward affine, which saves reading y in both backward passes. */
int kdl_bn_bwd(const void* dy, const void* x, const void* y, void* dx, void* dres, int64_t M,
               int C, const float* gamma, const float* beta, const float* save_mean_rstd, int relu,
               void* dgamma_bf16, void* dbeta_bf16, void* workspace, size_t workspace_bytes,
               cudaStream_t stream);

/* ReLU bit mask: kdl_bn_fwd_mask / kdl_bn_fwd_from_stats_mask additionally write one bit per element
 * ("positive before the ReLU", M*C/8 bytes, bit j of byte v = element 8v+j; may be NULL), and
 * kdl_bn_bwd_mask reads those bits instead of y.  For a ReLU that follows a residual add the mask cannot
 * be recomputed from x, and y is 16x the size of the bit mask. */
int kdl_bn_fwd_mask(const void* x, const void* res, void* y, int64_t M, int C, const float* gamma,
                    const float* beta, float* running_mean, float* running_var, float eps, float momentum,
                    int relu, float* save_mean_rstd, void* relu_mask, void* workspace, size_t workspace_bytes,
                    cudaStream_t stream);
int kdl_bn_bwd_mask(const void* dy, const void* x, const void* relu_mask, void* dx, void* dres, int64_t M, int C,
                    const float* gamma, const float* save_mean_rstd, void* dgamma_bf16, void* dbeta_bf16,
                    void* workspace, size_t workspace_bytes, cudaStream_t stream);
int kdl_bn_fwd_from_stats_mask(const void* x, const void* res, void* y, int64_t M, int C, const float* gamma,
                               const float* beta, float* running_mean, float* running_var, float eps,
                               float momentum, int relu, float* save_mean_rstd, const float* stats_partial,
                               int nparts, void* relu_mask, void* workspace, size_t workspace_bytes,
                               cudaStream_t stream);

/* Finalisation only (mean / rstd, running statistics, forward coefficients coef_out = [2][C]) from epilogue
 * partials, and a 3x3/s2 max-pool that applies relu(a*x + b) on load: BatchNorm apply + ReLU + max-pool of
 * the ResNet stem in one pass, the normalised activation is never materialised. */
int kdl_bn_finalize_from_stats(int64_t M, int C, const float* gamma, const float* beta, float* running_mean,
                               float* running_var, float eps, float momentum, float* save_mean_rstd,
                               const float* stats_partial, int nparts, float* coef_out, cudaStream_t stream);
int kdl_maxpool3x3s2_affine_relu_fwd(const void* x, const float* coef, void* y, void* idx_u8, int N, int H, int W,
                                     int C, cudaStream_t stream);

/* BN forward fed by the column statistics a preceding kdl_gemm_bf16_ex call produced in its
 * epilogue ([nparts][2][C] partial sums): the separate statistics pass over x is skipped. */
int kdl_bn_fwd_from_stats(const void* x, const void* res, void* y, int64_t M, int C, const float* gamma,
                          const float* beta, float* running_mean, float* running_var, float eps,
                          float momentum, int relu, float* save_mean_rstd, const float* stats_partial,
                          int nparts, void* workspace, size_t workspace_bytes, cudaStream_t stream);

/* kdl_bn_fwd for callers that cannot allocate device memory (the GPU's memory is wholly owned by
 * someone else's pool, or cudaMalloc's device-wide synchronisation is not acceptable where the
 * call is made).  This entry point has no counterpart in the reference.  Same arguments as
 * kdl_bn_fwd minus the workspace: the scratch is a fixed KDL_BN_STATIC_SCRATCH_BYTES array built
 * into the library (held on every GPU that loads it, until exit).  Shapes with
 * kdl_bn_workspace_bytes(M, C) above that limit are rejected.  The first call on a device
 * resolves the array's address: make it before any stream capture.  Calls on one GPU share the
 * array, so calls that overlap on the GPU race; run them one after another. */
#define KDL_BN_STATIC_SCRATCH_BYTES (512 * 1024)
size_t kdl_bn_static_scratch_limit(void);
int kdl_bn_fwd_static_scratch(const void* x, const void* res, void* y, int64_t M, int C,
                              const float* gamma, const float* beta, float* running_mean,
                              float* running_var, float eps, float momentum, int relu,
                              float* save_mean_rstd, cudaStream_t stream);

/* ---- N8: other memory-bound fused ops of the DDP worker --------------------------------- */
int kdl_u8_normalize(const void* in_u8, void* out_bf16, int64_t npix, int cpad,
                     const float* mean3_host, const float* std3_host, cudaStream_t stream);
/* uint8 NHWC image -> normalised bf16 im2col patches [N*OH*OW, KP] of the 7x7/s2/p3 stem
 * convolution (KP = 168: column = kh*24 + kw*3 + c, columns 21..23 of every filter row zero), so the
 * stem runs on kdl_gemm_bf16.  W % 4 == 0, W <= 256. */
int kdl_stem_im2col_u8(const void* img_u8, void* patches_bf16, int N, int H, int W, int KP,
                       const float* mean3_host, const float* std3_host, cudaStream_t stream);
size_t kdl_softmax_xent_workspace_bytes(int B);
int kdl_softmax_xent(const void* logits, int logits_is_fp32, const int64_t* labels,
                     void* dlogits_bf16, float* loss_out, int B, int V, int ldv, float grad_scale,
                     void* workspace, size_t workspace_bytes, cudaStream_t stream);
int kdl_sgd_momentum(float* p32, float* mom, const void* g16, void* p16, int64_t n,
                     int64_t wd_count, float lr, float mu, float wd, float grad_scale,
                     cudaStream_t stream);
int kdl_add_bf16(const void* a, const void* b, void* out, int64_t n, cudaStream_t stream);
int kdl_maxpool3x3s2_fwd(const void* x, void* y, void* idx_u8, int N, int H, int W, int C,
                         cudaStream_t stream);
int kdl_maxpool3x3s2_bwd(const void* dy, const void* idx_u8, void* dx, int N, int H, int W, int C,
                         cudaStream_t stream);
int kdl_avgpool_fwd(const void* x, void* y, int N, int HW, int C, cudaStream_t stream);
int kdl_avgpool_bwd(const void* dy, void* dx, int N, int HW, int C, cudaStream_t stream);

/* ---- N6: bf16 GEMM on tcgen05 tensor cores (TMEM accumulators, TMA loads/stores) -----------
 * C[M,N] = op(A) op(B) (+bias[N]) (+addend[M,N]) (ReLU).  a_mn_major=0: A is [M, lda] (K
 * contiguous); 1: A is [K, lda] (M contiguous).  Same for B with N.  nt = forward of a 1x1
 * convolution / linear layer, nn = dgrad, tn = wgrad (split over K; `scratch` of
 * kdl_gemm_scratch_bytes() holds the fp32 partial tiles, reduced in a fixed order).  Replaces the
 * cuBLAS/cuDNN calls of the stock worker image (reference
 * example/pytorch/pytorch_job_mnist_mpi.yaml:15). */
int kdl_gemm_splits(int M, int N, int K, int requested);
size_t kdl_gemm_scratch_bytes(int M, int N, int K, int split_k);
int kdl_gemm_bf16(const void* A, const void* B, void* C, int M, int N, int K, int a_mn_major,
                  int b_mn_major, int lda, int ldb, int ldc, const void* bias_bf16,
                  const void* addend_bf16, int relu, int split_k, void* scratch, size_t scratch_bytes,
                  cudaStream_t stream);
/* Same, and additionally (stats_partial != NULL) column sum / sum-of-squares of the bf16 output, one
 * partial row per CTA: [kdl_gemm_max_ctas()][2][N] floats (kdl_gemm_max_ctas() = the most rows
 * any launch writes); *stats_parts_out = number of rows written. */
int kdl_gemm_max_ctas(void);
int kdl_gemm_bf16_ex(const void* A, const void* B, void* C, int M, int N, int K, int a_mn_major,
                     int b_mn_major, int lda, int ldb, int ldc, const void* bias_bf16,
                     const void* addend_bf16, int relu, int split_k, void* scratch, size_t scratch_bytes,
                     float* stats_partial, int* stats_parts_out, cudaStream_t stream);

/* C[M, N] = A B + up-sampled addend: the rows of C are the pixels of an [n, H, W] map and `addend_compact`
 * is [n, H/2, W/2, N] (row stride ldc), added at even (h, w) only.  Used for the input gradient of a
 * bottleneck whose 1x1 / stride-2 downsample branch reads the same input: that branch's dgrad stays a
 * compact GEMM result and is never scattered into a mostly-zero full-size tensor. */
int kdl_gemm_bf16_subadd(const void* A, const void* B, void* C, int M, int N, int K, int a_mn_major,
                         int b_mn_major, int lda, int ldb, int ldc, const void* addend_compact, int H, int W,
                         cudaStream_t stream);
/* C = A B + (mask ? addend : 0): relu_mask is the [M*N/8] bit mask of kdl_bn_fwd_mask (bit j of byte v <->
 * element 8v+j).  N % 64 == 0, ldc == N.  The stock worker materialises dOut * mask in a separate
 * elementwise kernel (ATen threshold_backward); the reference operator has no math (reference
 * example/pytorch/pytorch_job_mnist_mpi.yaml:15 names an image). */
int kdl_gemm_bf16_maskadd(const void* A, const void* B, void* C, int M, int N, int K, int a_mn_major,
                          int b_mn_major, int lda, int ldb, int ldc, const void* addend,
                          const void* relu_mask, cudaStream_t stream);

/* ---- N7c: the 7x7 / stride-2 stem convolution straight from the uint8 image (224x224) -----------------
 * Builder warps normalise the image rows and write the patch tile of one output row directly in the
 * swizzled shared-memory layout tcgen05.mma reads; no patch matrix exists in global memory.  w / gw are
 * [64][168] in the kh*24 + kw*3 + c column layout.  fwd optionally emits BatchNorm column statistics
 * ([kdl_stem_conv_max_parts()][2][64]); wgrad keeps D[64,192] in TMEM across all rows of a CTA and needs
 * kdl_stem_conv_wgrad_scratch_bytes() of scratch.  No reference counterpart. */
int kdl_stem_conv_supported(int H, int W);
int kdl_stem_conv_max_parts(void);
size_t kdl_stem_conv_wgrad_scratch_bytes(void);
int kdl_stem_conv_fwd(const void* img_u8, const void* w_padded, void* y, int N, int H, int W, const float* mean3,
                      const float* std3, float* stats_partial, int* stats_parts_out, cudaStream_t stream);
int kdl_stem_conv_wgrad(const void* img_u8, const void* dy, void* gw_padded, int N, int H, int W,
                        const float* mean3, const float* std3, void* scratch, size_t scratch_bytes,
                        cudaStream_t stream);

/* ---- N7d: the stem as a 4x4 / stride-1 convolution on a 2x2 space-to-depth image (224x224) ------------
 * kdl_stem_s2d_pack writes S [N][115][115][16] bf16 (normalised, zero-bordered; kdl_stem_s2d_image_bytes(N)) once per
 * step; fwd and wgrad then read their operand tiles with TMA through a tensor map whose 32-byte pixel stride is
 * smaller than its 128-byte inner extent (overlapping 4-pixel windows) — no patch building by threads.  w / gw are
 * [64][168] in the kh*24 + kw*3 + c column layout as for N7c; w2_scratch is 64*256 bf16.  fwd optionally emits
 * BatchNorm column statistics ([kdl_stem_conv_max_parts()][2][64]); wgrad needs
 * kdl_stem_s2d_wgrad_scratch_bytes() of scratch.  No reference counterpart. */
size_t kdl_stem_s2d_image_bytes(int N);
size_t kdl_stem_s2d_wgrad_scratch_bytes(void);
int kdl_stem_s2d_pack(const void* img_u8, void* s2d, int N, int H, int W, const float* mean3, const float* std3,
                      cudaStream_t stream);
int kdl_stem_s2d_fwd(const void* s2d, const void* w_padded, void* w2_scratch, void* y, int N, int H, int W,
                     float* stats_partial, int* stats_parts_out, cudaStream_t stream);
int kdl_stem_s2d_wgrad(const void* s2d, const void* dy, void* gw_padded, int N, int H, int W, void* scratch,
                       size_t scratch_bytes, cudaStream_t stream);

/* ---- N7: implicit-GEMM convolution on the same tcgen05 kernel (NHWC bf16) ------------------
 * The activation operand is fetched with TMA im2col loads (no patch matrix in memory).  mode 0 =
 * fprop (x, w -> y), 1 = dgrad (dy, w -> dx [+addend]; stride-1 "same" convolutions), 2 = wgrad
 * (dy, x -> dw, split over the pixels, `scratch` of kdl_conv2d_scratch_bytes()).  Cin, Cout
 * multiples of 64.  Replaces the cuDNN convolutions of the stock worker. */
size_t kdl_conv2d_scratch_bytes(int mode, int N, int H, int W, int Cin, int Cout, int R, int S, int stride,
                                int pad);
int kdl_conv2d_bf16(int mode, const void* in0, const void* in1, void* out, int N, int H, int W, int Cin,
                    int Cout, int R, int S, int stride, int pad, const void* addend_bf16, void* scratch,
                    size_t scratch_bytes, cudaStream_t stream);
/* Input gradient of a 3x3 / stride-2 / pad-1 convolution: dy [N, H/2, W/2, Cout], w [Cout, 3, 3, Cin] ->
 * dx [N, H, W, Cin] (H, W even).  Four stride-1 sub-convolutions by input-pixel parity on the TMA-im2col
 * kernel (1 + 2 + 2 + 4 taps).  The stock worker gets this from cuDNN through ATen convolution_backward; the
 * reference operator has no math (reference example/pytorch/pytorch_job_mnist_mpi.yaml:15 names an image). */
int kdl_conv2d_dgrad_s2(const void* dy, const void* w, void* dx, int N, int H, int W, int Cin, int Cout,
                        cudaStream_t stream);
/* fprop (as kdl_conv2d_bf16 mode 0) that also returns the BatchNorm column statistics of y from its epilogue:
 * stats_partial [kdl_gemm_max_ctas()][2][Cout] fp32, *stats_parts_out valid rows (feed kdl_bn_fwd_from_stats).
 * The stock worker runs cuDNN's convolution and a separate batch-norm statistics kernel; no counterpart in the
 * reference operator (reference example/pytorch/pytorch_job_mnist_mpi.yaml:15 names an image). */
int kdl_conv2d_fprop_stats(const void* x, const void* w, void* y, int N, int H, int W, int Cin, int Cout, int R,
                           int S, int stride, int pad, float* stats_partial, int* stats_parts_out,
                           cudaStream_t stream);

/* ---- N7b: 3x3 / stride-1 / pad-1 convolution with the input halo resident in shared memory ----
 * One tiled TMA box per 64 channels brings a (rows+2) x (W+2) input halo into shared memory; the nine
 * filter taps are row-shifted UMMA descriptor windows over that one buffer, the CTA's weights stay
 * resident.  mode 0 = fprop, 1 = dgrad (optionally + addend, ReLU).  Supported when Cin, Cout are
 * multiples of 64, the weights of 64 output channels fit next to two halo buffers (C <= 128) and
 * some divisor TR of H has TR*(W+2) <= 128 — ask kdl_conv3x3_halo_supported().  With Cout (fprop) /
 * Cin (dgrad) of 64 or 128 a CTA pair runs one cta_group::2 MMA over two pixel tiles and all output
 * channels; dgrad then needs kdl_conv3x3_halo_scratch_bytes() of scratch for a flipped filter copy.  No reference
 * counterpart (the operator launches pods; cuDNN would run this in the stock worker). */
int kdl_conv3x3_halo_supported(int mode, int N, int H, int W, int Cin, int Cout);
size_t kdl_conv3x3_halo_scratch_bytes(int mode, int N, int H, int W, int Cin, int Cout);
int kdl_conv3x3_halo_bf16(int mode, const void* in0, const void* w, void* out, int N, int H, int W, int Cin,
                          int Cout, const void* addend_bf16, int relu, void* scratch, size_t scratch_bytes,
                          cudaStream_t stream);
/* wgrad of the same convolution (Cin, Cout <= 128): the pixel index is the MMA K dimension, every CTA
 * owns one filter row and keeps its 3 x Cin accumulator columns in TMEM across all of its pixel tiles;
 * fp32 partials in `scratch` (kdl_conv3x3_halo_wgrad_scratch_bytes(), 0 = shape unsupported) are reduced
 * in a fixed order by a second kernel. */
size_t kdl_conv3x3_halo_wgrad_scratch_bytes(int N, int H, int W, int Cin, int Cout);
int kdl_conv3x3_halo_wgrad_bf16(const void* dy, const void* x, void* dw, int N, int H, int W, int Cin, int Cout,
                                void* scratch, size_t scratch_bytes, cudaStream_t stream);

/* ---- N3: histogram GBDT, histogram build fused with the cross-GPU reduction ---------------
 * Replaces XGBoost gpu_hist + ncclAllReduce of histograms, for which the reference only injects
 * the tracker address (reference controllers/xgboost/pod.go:106-152).  peer_local / flag_ptrs are
 * HOST arrays of `world` device pointers into the symmetric heap (NULL when world == 1).   Shared-memory accumulators are 17-bit fixed point relative to gh_abs_max (bound on |g|, |h|;
 * 1 for logistic loss): integer shared-memory atomics are native, fp32 ones are CAS loops. */
int kdl_gbdt_grad_logistic(const float* pred, const float* label, void* gh, int64_t n,
                           cudaStream_t stream);
size_t kdl_gbdt_hist_bytes(int nodes, int F);
int kdl_gbdt_hist_allreduce(const void* bins, int fstride, int F, const void* gh, const int32_t* ridx,
                            const int32_t* seg, const int32_t* build, int nbuild, int64_t n_rows_hint,
                            void* local_hist, void* out_hist, void* node_done, void* const* peer_local,
                            void* const* flag_ptrs, const void* mc_local /* NVLS alias of local_hist or NULL */,
                            int world, int rank, int max_nodes, unsigned epoch,
                            void* status, unsigned long long timeout_ns, float gh_abs_max,
                            void* epoch_state /* optional 2 x uint32 on the device, zeroed once: the epoch then lives
                            there (the `epoch` argument is ignored) and the launch can be replayed from a CUDA graph */,
                            cudaStream_t stream);
/* Rows grouped by tree node with one counting-sort pass (node ids of a level take <= 256 values):
 * ridx[n] row ids node-major, seg[nodes+1] first slot per node, counts[nodes]. */
size_t kdl_gbdt_partition_workspace_bytes(int nodes);
int kdl_gbdt_partition(const int32_t* node_of_row, int64_t n, int nodes, int32_t* ridx, int32_t* seg,
                       int32_t* counts, void* workspace, size_t workspace_bytes, cudaStream_t stream);
/* Per-level glue (a round is launch-latency bound at 8 GPUs): the smaller child of every parent from the
 * (global) row counts; both children's histograms from the built one and the parent (subtraction trick);
 * leaf values of the deepest level. */
int kdl_gbdt_choose_build(const void* counts, int counts_are_float, int nodes, int32_t* build, cudaStream_t stream);
int kdl_gbdt_assemble_hist(const void* parent_hist, const void* built_hist, const int32_t* build, int nodes, int F,
                           void* out_hist, cudaStream_t stream);
int kdl_gbdt_leaf_values(const float* splits8, int nodes, float lambda, float lr, float* leaf, cudaStream_t stream);
int kdl_gbdt_split_eval(const void* hist, int nodes, int F, float lambda, float gamma,
                        float min_child_weight, float* out8, cudaStream_t stream);
int kdl_gbdt_apply_split(int32_t* node_of_row, const void* bins, int fstride, const float* splits8,
                         int64_t n, cudaStream_t stream);
int kdl_gbdt_update_pred(float* pred, const int32_t* node_of_row, const float* leaf_value, int64_t n,
                         cudaStream_t stream);

/* ---- N4/N5: sharded embedding gather (+DIN attention pooling) and scatter-add SGD update ----
 * over NVLink peer memory; replaces all_to_all_single + embedding_bag / index_add_ of a stock
 * PyTorch worker (XDL's PS exchange; reference controllers/xdl/xdljob_controller.go:191-217).
 * shard_ptrs: HOST array of `world` device pointers to the fp32 row shards [rows_per_rank, 64].
 * interleaved = 1: row id lives on rank id % world at local row id / world (skewed ids load every NVLink
 * port alike); 0: rank r owns ids [r * rows_per_rank, (r + 1) * rows_per_rank). */
int kdl_din_gather_attend(void* const* shard_ptrs, int world, int64_t rows_per_rank, int interleaved,
                          const int64_t* user, const int64_t* cand, const int64_t* hist, int B, int L,
                          int D, void* x_bf16, void* hsave_bf16, float* asave, cudaStream_t stream);
int kdl_din_scatter_update(void* const* shard_ptrs, int world, int64_t rows_per_rank, int interleaved,
                           const int64_t* user, const int64_t* cand, const int64_t* hist, int B, int L,
                           int D, const void* x_bf16, const void* dx_bf16, const void* hsave_bf16,
                           const float* asave, float lr, cudaStream_t stream);
/* Generic dispatch / combine over a sharded fp32 table of any row width D (multiple of 4): the two halves of an
 * embedding all-to-all as single kernels over NVLink peer memory.  kdl_emb_gather: out[i, :] = E[ids[i], :] (fp32 or
 * bf16 output; ids < 0 give zero rows).  kdl_emb_scatter_add: E[ids[i], :] += scale * grad[i, :] with vectorised
 * red.global.add (duplicates coalesce; ids < 0 are skipped) — with scale = -lr it is sparse SGD on the owner.
 * The stock worker does all_to_all_single(ids) + index_select + all_to_all_single(vectors), and the reverse with
 * index_add_; the reference only wires TASK_NAME / TASK_INDEX (reference controllers/xdl/xdljob_controller.go:191-217). */
int kdl_emb_gather(void* const* shard_ptrs, int world, int64_t rows_per_rank, int interleaved, int D,
                   const int64_t* ids, int64_t n, void* out, int out_bf16, cudaStream_t stream);
int kdl_emb_scatter_add(void* const* shard_ptrs, int world, int64_t rows_per_rank, int interleaved, int D,
                        const int64_t* ids, int64_t n, const void* grad, int grad_bf16, float scale,
                        cudaStream_t stream);

/* The elementwise glue of the DIN MLP, fused (the stock worker runs ~29 ATen kernels for it; no counterpart in the
 * reference, see above).  kdl_din_bce_grad: mean BCE-with-logits of column 0 of z [B, ldz] and its gradient as the
 * padded [B, 8] bf16 matrix (column 0 = (sigmoid(z) - y) / B).  kdl_relu_bwd_colsum: dh = (h > 0 ? dh : 0) in place
 * (h may be NULL: no mask) and dbias = column sums of dh, deterministic; workspace zeroed once. */
int kdl_din_bce_grad(const void* z_bf16, int ldz, const float* label, int B, float* loss, void* dz_bf16,
                     cudaStream_t stream);
size_t kdl_relu_bwd_colsum_workspace_bytes(int N);
int kdl_relu_bwd_colsum(void* dh_bf16, const void* h_bf16, int M, int N, void* dbias_bf16, void* workspace,
                        size_t workspace_bytes, cudaStream_t stream);

/* ---- N1/N2: fused gradient-bucket all-reduce over NVLink peer memory ---------------------
 * Replaces ncclAllReduce + div_ + cast of torch DDP, for which the reference only injects
 * MASTER_ADDR/MASTER_PORT/WORLD_SIZE/RANK (reference
 * controllers/pytorch/pytorchjob_controller.go:180-233).  in/out/flag pointer tables are HOST
 * arrays of `world` device pointers (peer mappings of the same symmetric offset); flags must be
 * kdl_allreduce_flags_bytes(channels) bytes per rank, zero-initialised once.  `epoch_state`
 * is LOCAL device memory of the same size, zero-initialised once; the kernels keep the barrier
 * epochs there, so launches can be captured into a CUDA graph and replayed.  Calls on one
 * channel must be issued in the same order on all ranks.  mc_in/mc_out: optional NVLS
 * multicast aliases (NULL -> P2P path).  emulate=1 runs all ranks inside one cooperative
 * launch on one GPU (protocol test). */
size_t kdl_allreduce_flags_bytes(int channels);
int kdl_allreduce_fused(const void* const* in_ptrs, void* const* out_ptrs, void* const* flag_ptrs,
                        const void* mc_in, void* mc_out, int world, int rank, int64_t n,
                        int in_dtype, int out_dtype, float scale, void* epoch_state, int channels,
                        int channel, int max_blocks, int emulate, void* status, unsigned long long timeout_ns,
                        cudaStream_t stream);
/* Flag-only barrier over the same flag / epoch arrays (own `channel`, same call order on all ranks):
 * everything the ranks enqueued on `stream` before it has completed on EVERY rank before anything
 * enqueued after it starts on any rank.  Graph-replayable like the all-reduce.  No counterpart in
 * the reference (it has no communication backend; reference pkg/job_controller/service.go:237-295). */
int kdl_symm_barrier(void* const* flag_ptrs, int world, int rank, void* epoch_state, int channels,
                     int channel, int emulate, void* status, unsigned long long timeout_ns,
                     cudaStream_t stream);

/* ---- symmetric memory plumbing ------------------------------------------
 * Legacy CUDA IPC path (cudaMalloc + cudaIpc handles) and VMM path (cuMemCreate + POSIX fd
 * export, needed for NVLS multicast).  The reference has no communication backend at all; it
 * only provides DNS names and ports (reference pkg/job_controller/service.go:237-295). */
int kdl_symm_malloc(size_t bytes, void** ptr);
int kdl_symm_free(void* ptr);
int kdl_ipc_get_handle(void* ptr, void* handle64_out);
int kdl_ipc_open_handle(const void* handle64, void** ptr_out);
int kdl_ipc_close_handle(void* ptr);
int kdl_enable_peer_access(int peer_device);
int kdl_vmm_supported(int device, int* posix_fd_ok, int* multicast_ok);
size_t kdl_vmm_granularity(int device, int multicast_world);
int kdl_vmm_alloc(int device, size_t bytes, unsigned long long* handle_out, void** ptr_out,
                  int* fd_out);
int kdl_vmm_import(int device, int fd, size_t bytes, unsigned long long* handle_out,
                   void** ptr_out);
int kdl_vmm_release(unsigned long long handle, void* ptr, size_t bytes);
int kdl_mc_create(int world, size_t bytes, unsigned long long* mc_out, int* fd_out);
int kdl_mc_import(int fd, unsigned long long* mc_out);
int kdl_mc_add_device(unsigned long long mc, int device);
int kdl_mc_bind_and_map(unsigned long long mc, unsigned long long mem_handle, int device,
                        size_t bytes, void** mc_ptr_out);
int kdl_mc_release(unsigned long long mc, void* mc_ptr, size_t bytes);

#ifdef __cplusplus
}
#endif
#endif /* KUBEDL_B200_H_ */
