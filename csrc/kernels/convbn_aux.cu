// Small fp32 kernels around the folded 1x1-convolution + batch-norm GEMMs (convbn_tcgen05.cu).
//
// With X [M, K] the layer input, W [N, K] the kernel, y = X W^T never stored:
//   forward  finalize: Gram = X^T X and S = colsum(X) give, per output channel c,
//              mean_c = w_c . S / M,   E[y_c^2] = w_c^T Gram w_c / M
//            -> scale / shift for the GEMM epilogue, saved mean / inv-std, running statistics,
//               and U = W Gram (kept for the backward pass);
//   backward finalize: with P = g^T X and sum_g = colsum(g) (g = upstream gradient after the ReLU
//            mask) the batch-norm reductions are
//              sum(g * xhat)_c = rstd_c (sum_k P[c,k] W[c,k] - mean_c sum_g_c)
//            and with a = sum_g / M, b = sum(g xhat) / M, s' = gamma rstd, t = s' b rstd:
//              dW[c,k] = s'_c (P[c,k] - a_c S[k] - b_c rstd_c (U[c,k] - mean_c S[k]))
//              W'[c,k] = s'_c W[c,k]                  (first B operand of the input-gradient GEMM)
//              Q[j,k]  = sum_c t_c W[c,j] W[c,k]      (second B operand, stored negated)
//              bias[k] = - sum_c (s'_c a_c - t_c mean_c) W[c,k]
//            so that dX = g W' - X Q + bias.
// These are O(N K^2) flops on matrices of at most a few MB: plain CUDA-core kernels, one launch each.
//
// No counterpart in the reference (see convbn_tcgen05.cu).
#include "common.cuh"
#include "../../include/cloud_b200_kernels.h"

namespace cb2 {

constexpr int kAuxCopies = 8;        // replicas in the column-sum scratch (bn.cu / GEMM epilogues)
constexpr int kCH = 8;               // output channels per block

__device__ __forceinline__ float block_sum(float v, float* red /*[32]*/) {
  v = warp_sum(v);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __syncthreads();
  if (lane == 0) red[warp] = v;
  __syncthreads();
  const int nw = (blockDim.x + 31) >> 5;
  float t = 0.f;
  for (int w = 0; w < nw; ++w) t += red[w];
  return t;
}

// grid = N / kCH, block = K threads (thread j owns input channel j)
__global__ void __launch_bounds__(1024)
convbn_fwd_finalize_kernel(const float* __restrict__ gram, int ldg, const float* __restrict__ xsums,
                           const __nv_bfloat16* __restrict__ W, int ldw, const float* __restrict__ gamma,
                           const float* __restrict__ beta, float* __restrict__ run_mean,
                           float* __restrict__ run_var, const float* __restrict__ mean_bias,
                           float* __restrict__ scale, float* __restrict__ shift,
                           float* __restrict__ save_mean, float* __restrict__ save_rstd,
                           float* __restrict__ U, float* __restrict__ S_out, int N, int K, float M,
                           float eps, float momentum) {
  extern __shared__ float sm[];
  float* wsm = sm;                    // [kCH][K]
  float* red = sm + kCH * K;          // [32]
  const int j = threadIdx.x;
  const int c0 = blockIdx.x * kCH;
  float S = 0.f;
#pragma unroll
  for (int r = 0; r < kAuxCopies; ++r) S += xsums[static_cast<size_t>(r) * 2 * K + j];
  if (blockIdx.x == 0) S_out[j] = S;
#pragma unroll
  for (int ch = 0; ch < kCH; ++ch)
    wsm[ch * K + j] = __bfloat162float(W[static_cast<size_t>(c0 + ch) * ldw + j]);
  __syncthreads();
  float u[kCH];
#pragma unroll
  for (int ch = 0; ch < kCH; ++ch) u[ch] = 0.f;
  for (int k = 0; k < K; ++k) {
    const float g = gram[static_cast<size_t>(k) * ldg + j];     // symmetric: row k, column j
#pragma unroll
    for (int ch = 0; ch < kCH; ++ch) u[ch] = fmaf(wsm[ch * K + k], g, u[ch]);
  }
  const float invM = 1.f / M;
#pragma unroll
  for (int ch = 0; ch < kCH; ++ch) {
    const int c = c0 + ch;
    U[static_cast<size_t>(c) * K + j] = u[ch];
    const float w = wsm[ch * K + j];
    const float mean = block_sum(w * S, red) * invM;
    const float ey2 = block_sum(w * u[ch], red) * invM;
    if (j == 0) {
      const float var = fmaxf(ey2 - mean * mean, 0.f);
      const float rstd = rsqrtf(var + eps);
      const float sc = gamma[c] * rstd;
      scale[c] = sc;
      shift[c] = beta[c] - mean * sc;
      save_mean[c] = mean;
      save_rstd[c] = rstd;
      const float unbiased = M > 1.f ? var * M / (M - 1.f) : var;
      run_mean[c] = (1.f - momentum) * run_mean[c] + momentum * (mean + (mean_bias ? mean_bias[c] : 0.f));
      run_var[c] = (1.f - momentum) * run_var[c] + momentum * unbiased;
    }
  }
}

// grid = N / kCH, block = K threads
__global__ void __launch_bounds__(1024)
convbn_bwd_finalize_a_kernel(const float* __restrict__ P, int ldp, const float* __restrict__ gsums,
                             const __nv_bfloat16* __restrict__ W, int ldw, const float* __restrict__ U,
                             const float* __restrict__ S, const float* __restrict__ save_mean,
                             const float* __restrict__ save_rstd, const float* __restrict__ gamma,
                             float* __restrict__ dgamma, float* __restrict__ dbeta,
                             __nv_bfloat16* __restrict__ dW, int lddw, __nv_bfloat16* __restrict__ Ws,
                             float* __restrict__ coef /*[N][2]*/, int N, int K, float M) {
  __shared__ float red[32];
  const int k = threadIdx.x;
  const float Sk = S[k];
  const float invM = 1.f / M;
#pragma unroll 1
  for (int ch = 0; ch < kCH; ++ch) {
    const int c = blockIdx.x * kCH + ch;
    const float w = __bfloat162float(W[static_cast<size_t>(c) * ldw + k]);
    const float pk = P[static_cast<size_t>(c) * ldp + k];
    const float sum_gy = block_sum(pk * w, red);
    float sum_g = 0.f;
#pragma unroll
    for (int r = 0; r < kAuxCopies; ++r) sum_g += gsums[static_cast<size_t>(r) * 2 * N + c];
    const float mean = save_mean[c], rstd = save_rstd[c];
    const float sum_gx = rstd * (sum_gy - mean * sum_g);
    const float a = sum_g * invM, b = sum_gx * invM;
    const float sp = gamma[c] * rstd;
    const float t = sp * b * rstd;
    if (k == 0) {
      dgamma[c] += sum_gx;
      dbeta[c] += sum_g;
      coef[2 * c] = t;
      coef[2 * c + 1] = sp * a - t * mean;
    }
    const float u = U[static_cast<size_t>(c) * K + k];
    const float g = sp * (pk - a * Sk - b * rstd * (u - mean * Sk));
    __nv_bfloat16* dst = dW + static_cast<size_t>(c) * lddw + k;
    *dst = __float2bfloat16(__bfloat162float(*dst) + g);
    Ws[static_cast<size_t>(c) * K + k] = __float2bfloat16(sp * w);
  }
}

// grid = K / kCH, block = K threads: rows j0..j0+7 of Qn = -(W^T diag(t) W); block 0 also the bias
__global__ void __launch_bounds__(1024)
convbn_bwd_finalize_b_kernel(const __nv_bfloat16* __restrict__ W, int ldw, const float* __restrict__ coef,
                             __nv_bfloat16* __restrict__ Qn, float* __restrict__ nbias, int N, int K) {
  extern __shared__ float sm[];       // [kCH] row values + t + const per channel chunk
  const int k = threadIdx.x;
  const int j0 = blockIdx.x * kCH;
  float acc[kCH];
#pragma unroll
  for (int i = 0; i < kCH; ++i) acc[i] = 0.f;
  float bias = 0.f;
  // channels in chunks of 32: their (t, const) and their weights at rows j0.. go through smem
  float* tc = sm;                     // [32][2]
  float* wj = sm + 64;                // [32][kCH]
  for (int cb = 0; cb < N; cb += 32) {
    __syncthreads();
    if (k < 64) tc[k] = coef[2 * cb + k];
    for (int idx = k; idx < 32 * kCH; idx += blockDim.x) {
      const int cc = idx / kCH, i = idx % kCH;
      wj[idx] = __bfloat162float(W[static_cast<size_t>(cb + cc) * ldw + j0 + i]);
    }
    __syncthreads();
#pragma unroll 4
    for (int cc = 0; cc < 32; ++cc) {
      const float w = __bfloat162float(W[static_cast<size_t>(cb + cc) * ldw + k]);
      const float tw = tc[2 * cc] * w;
      bias = fmaf(tc[2 * cc + 1], w, bias);
#pragma unroll
      for (int i = 0; i < kCH; ++i) acc[i] = fmaf(tw, wj[cc * kCH + i], acc[i]);
    }
  }
#pragma unroll
  for (int i = 0; i < kCH; ++i)
    Qn[static_cast<size_t>(j0 + i) * K + k] = __float2bfloat16(-acc[i]);
  if (blockIdx.x == 0) nbias[k] = -bias;
}

}  // namespace cb2

using namespace cb2;

static int aux_shape_ok(int N, int K, const char* what) {
  if (N <= 0 || K < 64 || K > 1024 || (K & 63) || (N & 31)) {
    set_last_error("%s: unsupported shape N=%d K=%d (K a multiple of 64 in [64, 1024], N a multiple of 32)",
                   what, N, K);
    return CB2_ERR_UNSUPPORTED;
  }
  return 0;
}

extern "C" int cb2_convbn_fwd_finalize(const float* gram, int ldg, const float* xsums, const void* W,
                                       int ldw, const float* gamma, const float* beta, float* run_mean,
                                       float* run_var, const float* mean_bias, float* scale, float* shift,
                                       float* save_mean, float* save_rstd, float* U, float* S, int N,
                                       int K, int M, float eps, float momentum, void* stream) {
  int rc = aux_shape_ok(N, K, "convbn_fwd_finalize");
  if (rc) return rc;
  const size_t smem = (static_cast<size_t>(kCH) * K + 32) * sizeof(float);
  static bool configured = false;
  if (!configured) {
    CB2_CUDA_OK(cudaFuncSetAttribute(convbn_fwd_finalize_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (kCH * 1024 + 32) * 4));
    configured = true;
  }
  convbn_fwd_finalize_kernel<<<N / kCH, K, smem, reinterpret_cast<cudaStream_t>(stream)>>>(
      gram, ldg, xsums, reinterpret_cast<const __nv_bfloat16*>(W), ldw, gamma, beta, run_mean, run_var,
      mean_bias, scale, shift, save_mean, save_rstd, U, S, N, K, static_cast<float>(M), eps, momentum);
  return check_cuda(cudaGetLastError(), "convbn_fwd_finalize launch");
}

extern "C" int cb2_convbn_bwd_finalize(const float* P, int ldp, const float* gsums, const void* W, int ldw,
                                       const float* U, const float* S, const float* save_mean,
                                       const float* save_rstd, const float* gamma, float* dgamma,
                                       float* dbeta, void* dW, int lddw, void* Ws, void* Qn, float* nbias,
                                       float* coef, int N, int K, int M, void* stream) {
  int rc = aux_shape_ok(N, K, "convbn_bwd_finalize");
  if (rc) return rc;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  convbn_bwd_finalize_a_kernel<<<N / kCH, K, 0, st>>>(
      P, ldp, gsums, reinterpret_cast<const __nv_bfloat16*>(W), ldw, U, S, save_mean, save_rstd, gamma,
      dgamma, dbeta, reinterpret_cast<__nv_bfloat16*>(dW), lddw, reinterpret_cast<__nv_bfloat16*>(Ws), coef,
      N, K, static_cast<float>(M));
  CB2_CUDA_OK(cudaGetLastError());
  convbn_bwd_finalize_b_kernel<<<K / kCH, K, (64 + 32 * kCH) * sizeof(float), st>>>(
      reinterpret_cast<const __nv_bfloat16*>(W), ldw, coef, reinterpret_cast<__nv_bfloat16*>(Qn), nbias, N, K);
  return check_cuda(cudaGetLastError(), "convbn_bwd_finalize launch");
}
