/* cloud_b200 kernel library - C ABI (libcloud_b200_kernels.so, sm_100a only).
 *
 * Every function returns 0 on success or a CB2_ERR_* code; cb2_last_error() gives the message of
 * the calling thread's last failure.  Pointers are device pointers unless noted; `stream` is a
 * cudaStream_t passed as void*.  Nothing here synchronises the device.
 *
 * The reference (marvinx97/cloud, TensorFlow Cloud) has no kernels of its own: each entry cites
 * the reference workload that makes TensorFlow run the equivalent op.
 */
#ifndef CLOUD_B200_KERNELS_H_
#define CLOUD_B200_KERNELS_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CB2_OK 0
#define CB2_ERR_ARG 1
#define CB2_ERR_CUDA 2
#define CB2_ERR_UNSUPPORTED 3

#define CB2_DTYPE_F32 0
#define CB2_DTYPE_BF16 1

const char* cb2_last_error(void);
int cb2_version(void);

/* ------------------------------------------------------------------------------------------
 * Dense GEMM on tcgen05/TMEM fed by TMA.  C[M,N] = op(A)[M,K] * op(B)[N,K]^T (+ epilogue).
 *   a_mn = 0: A stored row-major [M, lda>=K]  (K-major);  a_mn = 1: stored [K, lda>=M] (MN-major)
 *   b_mn = 0: B stored row-major [N, ldb>=K]  (K-major);  b_mn = 1: stored [K, ldb>=N] (MN-major)
 * A/B are bf16, 16-byte aligned, pitches multiples of 8 elements.  C is bf16 or fp32.
 * Reference workload: Dense layers of README.md:71-75 (K1/K2/K3).
 * ------------------------------------------------------------------------------------------ */
#define CB2_GEMM_RELU 1u      /* max(x, 0) after bias */
#define CB2_GEMM_OUT_F32 2u   /* C is fp32 (default bf16) */
#define CB2_GEMM_ACCUM 4u     /* C += result (fp32 C only) - wgrad straight into a grad arena */
#define CB2_GEMM_SOFTMAX 8u   /* row softmax after bias; needs N <= block_n */
int cb2_gemm_bf16(const void* A, const void* B, void* C, const float* bias /*nullable, [N]*/,
                  const void* relu_mask /*nullable bf16 [M,ldmask]: C=0 where mask<=0*/, int M,
                  int N, int K, int lda, int ldb, int ldc, int ldmask, int a_mn, int b_mn,
                  unsigned flags, int block_n /*0 = auto, else 64|128|256*/, void* stream);
/* Test hook: override the shared-memory matrix-descriptor fields (16-byte units). */
/* cb2_gemm_bf16 plus: split_k (1 = off, 0 = choose; > 1 needs OUT_F32|ACCUM: K slices are summed
 * into C by TMA reduce-adds, so C must hold zeros or a running sum) and colstats (nullable):
 * [cb2_gemm_colstat_copies()][2][N] fp32 accumulators that receive, per output column, the sum and
 * the sum of squares of the bf16-rounded results - the batch statistics of a following batch
 * normalisation, for free while the tile is still on chip (bf16 C, no bias). */
int cb2_gemm_bf16_ex(const void* A, const void* B, void* C, const float* bias, const void* relu_mask,
                     int M, int N, int K, int lda, int ldb, int ldc, int ldmask, int a_mn, int b_mn,
                     unsigned flags, int block_n, int split_k, float* colstats, void* stream);
int cb2_gemm_colstat_copies(void);
/* out_bf16[M,N] += bf16(op(A) @ op(B)^T) with the reduction cut into split_k slices (0 = choose).
 * The slices meet in `workspace` (fp32 [M, ld_ws], all zero on entry) through TMA reduce-adds; the
 * epilogue warp that completes a 32-row slice rounds it into out_bf16, re-zeroes that part of the
 * workspace and resets its counter (`counters`: >= 4 * ceil(M/128) * ceil(N/block_n) unsigned, all
 * zero on entry), so both are ready for the next call without a memset.  Calls that share a
 * workspace / counters must be ordered on one stream.  Workload: 1x1-convolution weight gradients,
 * dW[Cout,Cin] = dY^T X over N*H*W rows, written straight into the bf16 gradient arena. */
int cb2_gemm_splitk_bf16(const void* A, const void* B, float* workspace, void* out_bf16,
                         unsigned* counters, int n_counters, int M, int N, int K, int lda, int ldb,
                         int ld_ws, int ld_out, int a_mn, int b_mn, int block_n, int split_k,
                         void* stream);
/* ResNet stem (7x7 / stride 2 / pad 3, 3 -> 64 channels) as an implicit GEMM on the tcgen05 kernel.
 * xs: bf16 [B, OH+3, OW+3, 16] space-to-depth image (pixel (r, s) = input pixels (2r-4+a, 2s-4+b),
 * channel (a*2+b)*3 + c, zero where the input has none); wm: bf16 [64, 256] kernel with
 * k = i*64 + j*16 + (a*2+b)*3 + c for kernel row 2i+a-1, column 2j+b-1; y / dy: bf16 [B, OH, OW, 64];
 * dwt: fp32 [256, 64], the transposed gradient of wm, accumulated (zero or a running sum on entry).
 * OW <= 128.  No counterpart in the reference (TensorFlow hands this convolution to cuDNN). */
int cb2_conv_stem_forward(const void* xs, const void* wm, void* y, int B, int OH, int OW, void* stream);
int cb2_conv_stem_wgrad(const void* xs, const void* dy, float* dwt, int B, int OH, int OW, void* stream);
/* Weight gradient of a k x k convolution with pad = k/2 and stride 1 or 2, bf16 NHWC
 * x [B, H*stride, W*stride, Cin] and dy [B, H, W, Cout] read in place through 4-D TMA maps (taps =
 * shifted pixel boxes, stride = TMA traversal stride, zero padding = TMA out-of-bounds fill):
 * dwt[(i*k + j)*Cin + ci, co] += sum_pixels x[.., h*stride+i-pad, w*stride+j-pad, ci] * dy[.., h, w, co];
 * dwt fp32 [k*k*Cin, Cout], zero or a running sum on entry.  W <= 64, Cin and Cout multiples of 64.
 * No counterpart in the reference (cuDNN inside TensorFlow). */
int cb2_conv_wgrad(const void* x, const void* dy, float* dwt, int B, int H, int W, int Cin, int Cout,
                   int ksize, int pad, int stride, void* stream);
/* 3x3 / stride 1 / pad 1 convolution as an implicit GEMM (csrc/kernels/conv3x3_tcgen05.cu), bf16 NHWC:
 *   input_gradient = 0: y[B,H,W,Cout] = conv(x[B,H,W,Cin], w[Cout,3,3,Cin])
 *   input_gradient = 1: "x" is dy [B,H,W,Cout], "y" is dx [B,H,W,Cin], w the same tensor (taps mirrored,
 *                       weight tiles read channel-transposed: no transposed copy)
 * One TMA load per 64 input channels brings a strip of the zero-padded image (rows of W + 2 pixels) into
 * shared memory; the nine taps are nine operand descriptors into that strip.  colstats (nullable) =
 * [8][2][channels produced] fp32, accumulated: per-channel sums / sums of squares of the stored values
 * over the real pixels - the batch statistics of a following batch-norm (cb2_bn_forward takes them).
 * 6 <= W <= 62, Cin and Cout multiples of 64, dense 16-byte aligned tensors; cb2_conv3x3_supported says
 * whether a geometry is accepted.  Workload: conv*_block*_2_conv of Keras' ResNet-50
 * (core/tests/examples/call_run_within_script_with_keras_fit.py:79-83); no counterpart in the
 * reference, which leaves it to cuDNN inside TensorFlow. */
int cb2_conv3x3_supported(int H, int W, int Cin, int Cout);
int cb2_conv3x3(const void* x, const void* w, void* y, int B, int H, int W, int Cin, int Cout,
                int input_gradient, float* colstats, void* stream);
/* Diagnostics / tuning only (bench/conv3_probe.py): descriptor base-offset mode, pipeline depths,
 * staging buffers, traffic to skip (0 everywhere = defaults). */
int cb2_conv3x3_debug(int base_off, int a_stages, int b_stages, int out_bufs, int skip);
/* Diagnostics only (bench/gemm_epilogue_probe.py): bit mask of epilogue phases to skip. */
int cb2_gemm_debug_epilogue(int bits);
int cb2_gemm_debug_descriptors(int enable, unsigned a_lbo, unsigned a_sbo, unsigned b_lbo,
                               unsigned b_sbo, unsigned a_kadv, unsigned b_kadv);

/* ------------------------------------------------------------------------------------------
 * Fused sparse softmax cross-entropy, forward + backward in one pass (K4, K11).
 *   dlogits = (softmax(logits) - onehot(label)) * grad_scale        (same dtype as logits)
 *   stats[0] += sum_b loss_b ; stats[1] += #(argmax == label) ; row_loss[b] = loss_b (nullable)
 * Reference workload: loss='sparse_categorical_crossentropy' README.md:77.
 * ------------------------------------------------------------------------------------------ */
int cb2_softmax_xent(const void* logits, int dtype, const int64_t* labels, void* dlogits,
                     float* row_loss, float* stats, int B, int C, int ld, float grad_scale,
                     void* stream);

/* cb2_softmax_xent plus: loss_out (nullable, one float, zero on entry) += this call's summed loss x
 * grad_scale; add_count != 0: stats[2] += B (the sample counter next to loss and hit count). */
int cb2_softmax_xent_ex(const void* logits, int dtype, const int64_t* labels, void* dlogits,
                        float* row_loss, float* stats, int B, int C, int ld, float grad_scale,
                        float* loss_out, int add_count, void* stream);

/* dz = dy * (y > 0) (nullable y => dz = dy), db[n] (+)= sum_m dz[m,n].  bf16 tensors. (K3) */
int cb2_relu_bwd_bias_grad(const void* dy, const void* y, void* dz, void* db, int db_dtype, int M,
                           int N, int ld, int accumulate, void* stream);
/* The same with dz = scale * dy * (y > 0): the backward of Dense + ReLU + Dropout in one pass, y
 * being the DROPOUT output (zero where the unit was dropped or inactive) and scale = 1 / (1 - rate);
 * dz nullable (bias gradient only). */
int cb2_relu_bwd_bias_grad_ex(const void* dy, const void* y, void* dz, void* db, int db_dtype, int M,
                              int N, int ld, int accumulate, float scale, void* stream);

/* Philox dropout, mask recomputed from (seed, offset, *step_ptr) in backward:
 * y = x * keep / (1-p).  step_ptr (nullable) is a device float holding the training step, so a
 * CUDA-graph replay draws a new mask each step. (K5) */
int cb2_dropout(const void* x, void* y, size_t n, float p, uint64_t seed, uint64_t offset,
                const float* step_ptr, int dtype, void* stream);

/* uint8 NHWC image -> bf16 NHWC, (x/255 - mean[c]) / std[c].  Input pipeline tail (K12). */
int cb2_u8_to_bf16_normalize(const uint8_t* in, void* out, size_t n_pixels, int channels,
                             const float* mean3, const float* inv_std3, void* stream);
/* Same normalisation, written in the space-to-depth layout cb2_conv_stem_forward reads:
 * out bf16 [B, H/2+3, W/2+3, 16] (see there); in uint8 [B, H, W, 3], H and W even. */
int cb2_u8_to_s2d_bf16(const uint8_t* in, void* out, int B, int H, int W, const float* mean3,
                       const float* inv_std3, void* stream);

/* ------------------------------------------------------------------------------------------
 * Fused NHWC batch normalisation (+ residual add) (+ ReLU), bf16 activations [M = N*H*W, C],
 * fp32 statistics and parameters (K8).  C must be 8, 16, 32 or a multiple of 64, <= 2048.
 * `sums` is caller-provided scratch of cb2_bn_scratch_floats(C) floats that MUST be zero on entry
 * (the reductions accumulate into it with atomics).
 *   forward : y = act(gamma * (x - mean) / sqrt(var + eps) + beta (+ residual)); training != 0 uses
 *             batch statistics, refreshes run_mean/run_var (momentum = update rate; mean_bias,
 *             nullable, is added to the mean that goes into run_mean) and saves mean / inv-std.
 *             training == 2: `sums` already holds the batch sums (written by the colstats
 *             epilogue of cb2_gemm_bf16_ex) and the statistics pass is skipped.
 *   backward: dx, dbeta, dgamma (and dres = masked dy when a residual was added before the ReLU);
 *             y nullable: with relu != 0 and y == NULL the ReLU mask is recomputed from x.
 * Reference workload: the 53 BatchNormalization layers of ResNet-50
 * (core/tests/examples/call_run_within_script_with_keras_fit.py:79-83).
 * ------------------------------------------------------------------------------------------ */
int cb2_bn_supported(int C);
int cb2_bn_scratch_floats(int C);
int cb2_bn_forward(const void* x, const void* residual /*nullable*/, void* y, float* sums,
                   const float* gamma, const float* beta, float* run_mean, float* run_var,
                   float* save_mean, float* save_invstd, const float* mean_bias /*nullable*/, int M,
                   int C, float eps, float momentum, int relu, int training, void* stream);
int cb2_bn_backward(const void* dy, const void* y /*nullable*/, const void* x,
                    const float* save_mean, const float* save_invstd, const float* gamma,
                    const float* beta, float* sums, float* dbeta, float* dgamma, void* dx,
                    void* dres /*nullable*/, int M, int C, int relu, void* stream);
/* cb2_bn_backward with a second incoming gradient: dy_total = dy + dy2 (dy2 nullable), summed in
 * fp32 inside both passes - for a batch-norm output that feeds two consumers (a residual block's
 * input), instead of an elementwise add of two activation-sized tensors in front of it. */
int cb2_bn_backward2(const void* dy, const void* dy2, const void* y, const void* x,
                     const float* save_mean, const float* save_invstd, const float* gamma,
                     const float* beta, float* sums, float* dbeta, float* dgamma, void* dx, void* dres,
                     int M, int C, int relu, void* stream);
/* The shortcut variant (BN + residual + ReLU) without re-reading the output in backward:
 * cb2_bn_forward_mask also writes mask_out[M * C / 8], one bit per element (output > 0), and
 * cb2_bn_backward3 with y_is_mask != 0 takes those bytes in place of y - 1/16 of the bytes in each
 * of the two backward passes. */
int cb2_bn_forward_mask(const void* x, const void* residual, void* y, float* sums, const float* gamma,
                        const float* beta, float* run_mean, float* run_var, float* save_mean,
                        float* save_invstd, const float* mean_bias, void* mask_out, int M, int C,
                        float eps, float momentum, int relu, int training, void* stream);
int cb2_bn_backward3(const void* dy, const void* dy2, const void* y, int y_is_mask, const void* x,
                     const float* save_mean, const float* save_invstd, const float* gamma,
                     const float* beta, float* sums, float* dbeta, float* dgamma, void* dx, void* dres,
                     int M, int C, int relu, void* stream);

/* Column sums / sums of squares of a bf16 [M, C] matrix into `sums` (cb2_bn_scratch_floats(C) floats,
 * accumulated: zero on entry) - the statistics pass of cb2_bn_forward on its own. */
int cb2_bn_stats(const void* x, float* sums, int M, int C, void* stream);
/* g = (dy (+ dy2, nullable)) where the sign bit saved by cb2_bn_forward_mask / cb2_convbn_gemm is set
 * (mask nullable: all kept), else 0; sums[0:C] += column sums of g.  The masked gradient of a
 * BN + shortcut + ReLU output as a GEMM operand, for the folded backward below. */
int cb2_bn_mask_grad(const void* dy, const void* dy2, const void* mask, void* g, float* sums, int M, int C,
                     void* stream);

/* ------------------------------------------------------------------------------------------
 * 1x1 convolution + batch normalisation (+ shortcut) (+ ReLU) folded into GEMMs: the convolution
 * output y = X W^T (X bf16 [M, K], W bf16 [N, K]) is never stored.
 *   forward : Gram = X^T X and column sums of X (cb2_gemm_bf16_ex, cb2_bn_stats) ->
 *             cb2_convbn_fwd_finalize: mean_c = w_c . S / M, E[y_c^2] = w_c^T Gram w_c / M ->
 *             scale / shift, saved mean / inv-std, running statistics, U = W Gram (u_ready != 0: U
 *             was computed by the caller, e.g. on the tensor cores from cb2_split_hilo_bf16 halves) ->
 *             cb2_convbn_gemm: z = relu(y * scale + shift + residual), one sign bit per element.
 *   backward: g = masked upstream gradient (cb2_bn_mask_grad, or the epilogue of the GEMM that
 *             produced it: mask_in + residual + colstats), P = g^T X ->
 *             cb2_convbn_bwd_finalize: dgamma, dbeta, dW (added into a bf16 arena), Ws = s' W,
 *             Wt = -t W, nbias (accumulated: zero on entry) -> Qn = Wt^T W (cb2_gemm_bf16) ->
 *             cb2_convbn_gemm over two operand pairs: dX = g Ws + X Qn + nbias.
 * cb2_convbn_gemm: C[M,N] = mask_in ? relu?((A B^T [+ A2 B2^T]) * scale + shift + residual) : 0, bf16,
 * N / K1 / K2 multiples of 64; b_mn = 0: B stored [N, K], 1: stored [K, N]; scale / shift / residual /
 * masks / colstats nullable; masks are bit arrays [M, N/32] words (bit set = element kept / positive);
 * colstats = [cb2_convbn_stat_copies()][2][N] fp32 column sums / sums of squares of the stored values,
 * accumulated.  block_n 0 = choose.
 * Workload: conv*_block*_3_conv/_3_bn/_add/_out and *_0_conv/_0_bn of Keras' ResNet-50
 * (core/tests/examples/call_run_within_script_with_keras_fit.py:79-83); no counterpart in the
 * reference, which leaves these to cuDNN convolution, batch-norm, add and ReLU kernels.
 * ------------------------------------------------------------------------------------------ */
int cb2_convbn_stat_copies(void);
int cb2_convbn_gemm(const void* A, int lda, const void* B, int ldb, int b_mn, const void* A2, int lda2,
                    const void* B2, int ldb2, int K2, void* C, int ldc, int M, int N, int K1,
                    const float* scale, const float* shift, const void* residual, int ldr, int relu,
                    const void* mask_in, void* mask_out, float* colstats, int block_n, void* stream);
int cb2_convbn_fwd_finalize(const float* gram, int ldg, const float* xsums, const void* W, int ldw,
                            const float* gamma, const float* beta, float* run_mean, float* run_var,
                            const float* mean_bias, float* scale, float* shift, float* save_mean,
                            float* save_rstd, float* U, float* S, int N, int K, int M, float eps,
                            float momentum, int u_ready, void* stream);
int cb2_convbn_bwd_finalize(const float* P, int ldp, const float* gsums, const void* W, int ldw,
                            const float* U, const float* S, const float* save_mean, const float* save_rstd,
                            const float* gamma, float* dgamma, float* dbeta, void* dW, int lddw, void* Ws,
                            void* Wt, float* nbias, int N, int K, int M, void* stream);
/* hi[i] = bf16(x[i]), lo[i] = bf16(x[i] - hi[i]): an fp32 matrix as two bf16 GEMM operands. */
int cb2_split_hilo_bf16(const float* x, void* hi, void* lo, size_t n, void* stream);
/* Every s-th pixel of an NHWC bf16 tensor (backward = 0: y[B, ceil(H/s), ceil(W/s), C] <- x) and its
 * adjoint (backward = 1: x is the gradient of the subsampled tensor, y the full-size gradient,
 * written in one pass, zeros between the sampled pixels).  The input side of a strided 1x1
 * convolution (conv*_block1_{0,1}_conv of ResNet-50), which then runs as a plain GEMM. */
int cb2_subsample_nhwc(const void* x, void* y, int B, int H, int W, int C, int s, int backward, void* stream);

/* cb2_bn_forward for callers that cannot allocate device memory.  (No counterpart in the
 * reference: it has no kernels and no C ABI; this entry exists for embedders of this library.)
 *
 * Who needs it: a plug-in inside a serving process that gave all of HBM to its own pool at
 * start-up (a further cudaMalloc fails), a host that links the C ABI without torch and so has no
 * caching allocator to borrow scratch from, or a steady-state loop that cannot afford the
 * device-wide synchronisation of cudaMalloc / cudaFree.  Same arguments and results as
 * cb2_bn_forward minus `sums`: the scratch is a 128 KiB array inside the library image
 * (8 replicas x 2 x 2048 floats), which every GPU that loads the library keeps until the process
 * exits.  Limit: C <= cb2_bn_static_scratch_max_channels() (2048); larger C is rejected with
 * CB2_ERR_UNSUPPORTED and a message naming the limit.
 *
 * The call allocates nothing (no cudaMalloc / cudaFree, no allocating library routine) and
 * clears the part of the scratch it accumulates into before use, because the array still holds
 * the previous call's sums.  All calls on a device share that one array: two calls that may
 * overlap on the GPU (different streams, or the same stream from racing host threads) corrupt
 * each other's statistics - serialise them.  The first call on each device looks up the array's
 * device address (cudaGetSymbolAddress, with that device current) and caches it; make that first
 * call before starting any stream capture.  `training` is 0 (running statistics: the scratch is not
 * touched at all) or 1 (batch statistics); 2 - caller-reduced sums - has no meaning without a caller
 * scratch and is rejected.  `stream` must belong to the current device (checked). */
int cb2_bn_static_scratch_max_channels(void);
int cb2_bn_forward_static(const void* x, const void* residual /*nullable*/, void* y,
                          const float* gamma, const float* beta, float* run_mean, float* run_var,
                          float* save_mean, float* save_invstd, const float* mean_bias /*nullable*/,
                          int M, int C, float eps, float momentum, int relu, int training,
                          void* stream);

/* NHWC max pooling with implicit zero padding; forward keeps a 1-byte arg-max per output element,
 * backward gathers (no atomics).  bf16, C multiple of 8 (K8: pool1_pad + pool1_pool of ResNet). */
int cb2_maxpool_nhwc_forward(const void* x, void* y, void* argmax, int N, int H, int W, int C,
                             int OH, int OW, int k, int s, int pad, void* stream);
int cb2_maxpool_nhwc_backward(const void* dy, const void* argmax, void* dx, int N, int H, int W,
                              int C, int OH, int OW, int k, int s, int pad, void* stream);

/* ------------------------------------------------------------------------------------------
 * Optimizers over flat arenas (K6, K7).  `hyper` is a device float[16]:
 *   [0] lr [1] beta1 [2] beta2 [3] eps [4] weight_decay [5] step [6] 1-beta1^t [7] 1-beta2^t
 *   [8] grad_scale [9] momentum / rho [10] nesterov(0/1) [11] rmsprop momentum
 * cb2_optim_tick advances [5..7] on the device so the whole step can live in a CUDA graph.
 * Reference workload: optimizer=Adam README.md:78; SGD cloud_fit/tests/unit/client_test.py:87-89;
 * RMSprop testdata/keras_tuner_cifar_example.py:76.
 * ------------------------------------------------------------------------------------------ */
#define CB2_OPT_NONE 0        /* plain reduction: out = grad_scale * sum */
#define CB2_OPT_ADAM_KERAS 1  /* p -= lr*sqrt(bc2)/bc1 * m / (sqrt(v) + eps) */
#define CB2_OPT_ADAM_TORCH 2  /* p -= lr/bc1 * m / (sqrt(v/bc2) + eps) */
#define CB2_OPT_ADAMW 3       /* ADAM_TORCH with decoupled weight decay */
#define CB2_OPT_SGD 4         /* buf = mu*buf + g ; p -= lr*(nesterov ? g + mu*buf : buf) */
#define CB2_OPT_RMSPROP 5     /* ms = rho*ms+(1-rho)g^2 ; mom = mu*mom + lr*g/sqrt(ms+eps) ; p -= mom */
#define CB2_HYPER_LEN 16
int cb2_optim_tick(float* hyper, void* stream);
/* master fp32 [n] (nullable => param_out is the fp32 master), s1/s2 fp32 state (m,v | buf | ms,mom) */
int cb2_optim_flat(int opt, float* master, const void* grad, int grad_dtype, float* s1, float* s2,
                   void* param_out, int param_dtype, size_t n, const float* hyper, int zero_grad,
                   void* stream);

/* ------------------------------------------------------------------------------------------
 * Peer-memory collectives over NVLink/NVSwitch (K9, K10, K11), fused with the optimizer.
 * All ranks launch the same call with the same sizes, in the same order, on one stream each.
 * Reference: implied by MirroredStrategy()/MultiWorkerMirroredStrategy() core/preprocess.py:139-143.
 * ------------------------------------------------------------------------------------------ */
#define CB2_MAX_RANKS 8
#define CB2_COMM_MAX_BLOCKS 128
#define CB2_COMM_AUX_FLOATS 256
/* signal pad bytes each rank must provide (zero-initialised, symmetric): flags then aux floats */
#define CB2_COMM_PAD_BYTES (CB2_COMM_MAX_BLOCKS * CB2_MAX_RANKS * 4 + CB2_COMM_AUX_FLOATS * 4)

#define CB2_ALGO_ONESHOT 0  /* every rank reduces everything (replicated state); latency bound */
#define CB2_ALGO_TWOSHOT 1  /* reduce-scatter by peer loads -> update shard -> all-gather by peer stores */
#define CB2_ALGO_NVLS 2     /* multimem.ld_reduce (in-switch sum) -> update shard -> multimem.st */
#define CB2_ALGO_LL1 3      /* flag-in-payload pushes, no barrier: one-shot, replicated state (<= ~64 KiB) */
#define CB2_ALGO_LL2 4      /* flag-in-payload pushes, no barrier: two-shot, sharded state (<= a few MiB) */

typedef struct {
  int rank, world;
  void* grad_ptrs[CB2_MAX_RANKS];   /* this rank's mapping of every rank's gradient arena */
  void* param_ptrs[CB2_MAX_RANKS];  /* ... parameter / output arena */
  void* pad_ptrs[CB2_MAX_RANKS];    /* ... signal pad (CB2_COMM_PAD_BYTES) */
  void* mc_grad;                    /* multicast mapping of the gradient arena (NVLS) or NULL */
  void* mc_param;                   /* multicast mapping of the parameter arena or NULL */
  uint32_t* epochs;                 /* local uint32[CB2_COMM_MAX_BLOCKS], zero-initialised */
  void* ll_ptrs[CB2_MAX_RANKS];     /* this rank's mapping of every rank's LL receive buffer, or NULL */
  size_t ll_bytes;                  /* size of one LL buffer (zero-initialised, symmetric); a bucket of
                                       S bytes needs cb2_comm_ll_bytes(world, S) for LL1, S / world for LL2 */
  uint32_t* ll_state;               /* local uint32[2], zero-initialised, shared by every context that
                                       uses the same LL buffers */
} cb2_comm_ctx;
/* Bytes of LL buffer that let one slot hold `slot_payload` bytes of gradients. */
size_t cb2_comm_ll_bytes(int world, size_t slot_payload);

/* Reduce grad[off, off+len) over ranks, apply `opt` to the owned part, publish parameters.
 *   ONESHOT/LL1: state arrays cover the whole range, indexed [state_off + i].
 *   TWOSHOT/NVLS/LL2: this rank owns [off + rank*len/world, +len/world); state indexed
 *                 [state_off + i] for i in the owned shard.
 * len must be a multiple of 8*world, off a multiple of 8.  aux_len fp32 scalars living in the
 * signal pad are sum-reduced one-shot into aux_out (loss / accuracy counters) by block 0. */
int cb2_comm_fused(const cb2_comm_ctx* ctx, int algo, int opt, int grad_dtype, int param_dtype,
                   size_t off, size_t len, float* master, float* s1, float* s2, size_t state_off,
                   const float* hyper, int zero_grads, int aux_len, float* aux_out, int blocks,
                   void* stream);
/* Copy param[off, off+len) of `root` into every rank's parameter arena (initial variable sync). */
int cb2_comm_broadcast(const cb2_comm_ctx* ctx, int root, int dtype, size_t off, size_t len,
                       int use_multicast, int blocks, void* stream);
/* Device-side barrier across ranks (one block). */
int cb2_comm_barrier(const cb2_comm_ctx* ctx, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* CLOUD_B200_KERNELS_H_ */
