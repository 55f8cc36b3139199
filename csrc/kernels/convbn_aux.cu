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
//              Wt[c,k] = -t_c W[c,k]                  (-Q = Wt^T W, the second B operand: one more
//                                                       tensor-core GEMM over the N channels)
//              bias[k] = - sum_c (s'_c a_c - t_c mean_c) W[c,k]
//            so that dX = g W' - X Q + bias.
// One warp per output channel; the only O(N K^2) work left on CUDA cores is U = W Gram.
//
// No counterpart in the reference (see convbn_tcgen05.cu).
#include <algorithm>
#include <type_traits>

#include "common.cuh"
#include "../../include/cloud_b200_kernels.h"

namespace cb2 {

constexpr int kAuxCopies = 8;        // replicas in the column-sum scratch (bn.cu / GEMM epilogues)
constexpr int kWPB = 8;              // warps per block = output channels per block (one warp each)

// One warp per output channel c; lane l owns input channels j = l + 32 i (JPL of them).
//   u_j = sum_k W[c,k] Gram[k,j]   (Gram symmetric: row k is read, 128 contiguous bytes per i)
// grid = N / kWPB, block = 32 * kWPB, dynamic smem = (kWPB + 1) * K floats
template <int JPL>
__global__ void __launch_bounds__(32 * kWPB)
convbn_fwd_finalize_kernel(const float* __restrict__ gram, int ldg, const float* __restrict__ xsums,
                           const __nv_bfloat16* __restrict__ W, int ldw, const float* __restrict__ gamma,
                           const float* __restrict__ beta, float* __restrict__ run_mean,
                           float* __restrict__ run_var, const float* __restrict__ mean_bias,
                           float* __restrict__ scale, float* __restrict__ shift,
                           float* __restrict__ save_mean, float* __restrict__ save_rstd,
                           float* __restrict__ U, float* __restrict__ S_out, int N, int K, float M,
                           float eps, float momentum, int u_ready) {
  extern __shared__ float sm[];
  float* Ssm = sm;                    // [K]
  float* wsm = sm + K;                // [kWPB][K]
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int c = blockIdx.x * kWPB + warp;
  for (int j = threadIdx.x; j < K; j += blockDim.x) {
    float S = 0.f;
#pragma unroll
    for (int r = 0; r < kAuxCopies; ++r) S += xsums[static_cast<size_t>(r) * 2 * K + j];
    Ssm[j] = S;
    if (blockIdx.x == 0) S_out[j] = S;
  }
  for (int j = lane; j < K; j += 32)
    wsm[warp * K + j] = __bfloat162float(W[static_cast<size_t>(c) * ldw + j]);
  __syncthreads();
  float u[JPL];
#pragma unroll
  for (int i = 0; i < JPL; ++i) u[i] = 0.f;
  const float* wrow = wsm + warp * K;
  // eight Gram rows per step: their loads are all issued before the first FMA needs one (a warp
  // here has nothing else to hide the L2 latency behind)
  constexpr int KU = JPL <= 4 ? 8 : (JPL <= 8 ? 4 : 2);
  if (u_ready) {
    // U = W Gram came from the tensor cores (two bf16 GEMMs over a hi / lo split of Gram)
#pragma unroll
    for (int i = 0; i < JPL; ++i)
      if (lane + 32 * i < K) u[i] = U[static_cast<size_t>(c) * K + lane + 32 * i];
  }
  for (int k0 = 0; k0 < (u_ready ? 0 : K); k0 += KU) {
    float gv[KU][JPL];
#pragma unroll
    for (int kk = 0; kk < KU; ++kk) {
      const float* g = gram + static_cast<size_t>(k0 + kk) * ldg + lane;
#pragma unroll
      for (int i = 0; i < JPL; ++i) gv[kk][i] = (lane + 32 * i < K) ? __ldg(g + 32 * i) : 0.f;
    }
#pragma unroll
    for (int kk = 0; kk < KU; ++kk) {
      const float wk = wrow[k0 + kk];
#pragma unroll
      for (int i = 0; i < JPL; ++i) u[i] = fmaf(wk, gv[kk][i], u[i]);
    }
  }
  float ms = 0.f, es = 0.f;
#pragma unroll
  for (int i = 0; i < JPL; ++i) {
    const int j = lane + 32 * i;
    if (j < K) {
      if (!u_ready) U[static_cast<size_t>(c) * K + j] = u[i];
      ms = fmaf(wrow[j], Ssm[j], ms);
      es = fmaf(wrow[j], u[i], es);
    }
  }
  ms = warp_sum(ms);
  es = warp_sum(es);
  if (lane == 0) {
    const float invM = 1.f / M;
    const float mean = ms * invM;
    const float var = fmaxf(es * invM - mean * mean, 0.f);
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

// One warp per output channel.  Besides the batch-norm gradients and dW it prepares the operands of
// the input-gradient GEMM: Ws = s' W, Wt = -t W (so that Q = Wt^T W is one more GEMM), and the bias
// nbias[k] -= sum_c (s'_c a_c - t_c mean_c) W[c,k] (accumulated: the caller zeroes nbias).
// grid = N / kWPB, block = 32 * kWPB, dynamic smem = K floats
template <int JPL>
__global__ void __launch_bounds__(32 * kWPB)
convbn_bwd_finalize_kernel(const float* __restrict__ P, int ldp, const float* __restrict__ gsums,
                           const __nv_bfloat16* __restrict__ W, int ldw, const float* __restrict__ U,
                           const float* __restrict__ S, const float* __restrict__ save_mean,
                           const float* __restrict__ save_rstd, const float* __restrict__ gamma,
                           float* __restrict__ dgamma, float* __restrict__ dbeta,
                           __nv_bfloat16* __restrict__ dW, int lddw, __nv_bfloat16* __restrict__ Ws,
                           __nv_bfloat16* __restrict__ Wt, float* __restrict__ nbias, int N, int K,
                           float M) {
  extern __shared__ float sm[];       // [K] bias partial of this block
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int c = blockIdx.x * kWPB + warp;
  for (int j = threadIdx.x; j < K; j += blockDim.x) sm[j] = 0.f;
  __syncthreads();
  float w[JPL], pk[JPL];
  float sum_gy = 0.f;
#pragma unroll
  for (int i = 0; i < JPL; ++i) {
    const int j = lane + 32 * i;
    w[i] = pk[i] = 0.f;
    if (j < K) {
      w[i] = __bfloat162float(W[static_cast<size_t>(c) * ldw + j]);
      pk[i] = P[static_cast<size_t>(c) * ldp + j];
      sum_gy = fmaf(pk[i], w[i], sum_gy);
    }
  }
  sum_gy = warp_sum(sum_gy);
  float sum_g = 0.f;
#pragma unroll
  for (int r = 0; r < kAuxCopies; ++r) sum_g += gsums[static_cast<size_t>(r) * 2 * N + c];
  const float invM = 1.f / M;
  const float mean = save_mean[c], rstd = save_rstd[c];
  const float sum_gx = rstd * (sum_gy - mean * sum_g);
  const float a = sum_g * invM, b = sum_gx * invM;
  const float sp = gamma[c] * rstd;
  const float t = sp * b * rstd;
  const float cc = sp * a - t * mean;
  if (lane == 0) {
    dgamma[c] += sum_gx;
    dbeta[c] += sum_g;
  }
#pragma unroll
  for (int i = 0; i < JPL; ++i) {
    const int j = lane + 32 * i;
    if (j < K) {
      const float Sj = S[j];
      const float u = U[static_cast<size_t>(c) * K + j];
      const float g = sp * (pk[i] - a * Sj - b * rstd * (u - mean * Sj));
      __nv_bfloat16* dst = dW + static_cast<size_t>(c) * lddw + j;
      *dst = __float2bfloat16(__bfloat162float(*dst) + g);
      Ws[static_cast<size_t>(c) * K + j] = __float2bfloat16(sp * w[i]);
      Wt[static_cast<size_t>(c) * K + j] = __float2bfloat16(-t * w[i]);
      atomicAdd(&sm[j], cc * w[i]);
    }
  }
  __syncthreads();
  for (int j = threadIdx.x; j < K; j += blockDim.x) atomicAdd(nbias + j, -sm[j]);
}

}  // namespace cb2

using namespace cb2;

static int aux_shape_ok(int N, int K, const char* what) {
  if (N <= 0 || K < 64 || K > 1024 || (K & 63) || (N & 7)) {
    set_last_error("%s: unsupported shape N=%d K=%d (K a multiple of 64 in [64, 1024], N a multiple of 8)",
                   what, N, K);
    return CB2_ERR_UNSUPPORTED;
  }
  return 0;
}

// fp32 -> two bf16 arrays with hi + lo = x to ~16 mantissa bits: the operand form in which an fp32
// matrix goes through the bf16 tensor cores without losing what a variance needs
__global__ void split_hilo_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ hi,
                                  __nv_bfloat16* __restrict__ lo, size_t n) {
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const float v = x[i];
    const __nv_bfloat16 h = __float2bfloat16(v);
    hi[i] = h;
    lo[i] = __float2bfloat16(v - __bfloat162float(h));
  }
}

extern "C" int cb2_split_hilo_bf16(const float* x, void* hi, void* lo, size_t n, void* stream) {
  if (n == 0) return 0;
  const int blocks = static_cast<int>(std::min<size_t>((n + 255) / 256, 1184));
  split_hilo_kernel<<<blocks, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      x, reinterpret_cast<__nv_bfloat16*>(hi), reinterpret_cast<__nv_bfloat16*>(lo), n);
  return check_cuda(cudaGetLastError(), "split_hilo launch");
}

template <typename F>
static int dispatch_jpl(int K, F&& f) {
  const int per = (K + 31) / 32;
  if (per <= 2) return f(std::integral_constant<int, 2>());
  if (per <= 4) return f(std::integral_constant<int, 4>());
  if (per <= 8) return f(std::integral_constant<int, 8>());
  if (per <= 16) return f(std::integral_constant<int, 16>());
  return f(std::integral_constant<int, 32>());
}

extern "C" int cb2_convbn_fwd_finalize(const float* gram, int ldg, const float* xsums, const void* W,
                                       int ldw, const float* gamma, const float* beta, float* run_mean,
                                       float* run_var, const float* mean_bias, float* scale, float* shift,
                                       float* save_mean, float* save_rstd, float* U, float* S, int N,
                                       int K, int M, float eps, float momentum, int u_ready, void* stream) {
  int rc = aux_shape_ok(N, K, "convbn_fwd_finalize");
  if (rc) return rc;
  const size_t smem = static_cast<size_t>(kWPB + 1) * K * sizeof(float);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  return dispatch_jpl(K, [&](auto jpl) {
    constexpr int JPL = decltype(jpl)::value;
    convbn_fwd_finalize_kernel<JPL><<<N / kWPB, 32 * kWPB, smem, st>>>(
        gram, ldg, xsums, reinterpret_cast<const __nv_bfloat16*>(W), ldw, gamma, beta, run_mean, run_var,
        mean_bias, scale, shift, save_mean, save_rstd, U, S, N, K, static_cast<float>(M), eps, momentum, u_ready);
    return check_cuda(cudaGetLastError(), "convbn_fwd_finalize launch");
  });
}

extern "C" int cb2_convbn_bwd_finalize(const float* P, int ldp, const float* gsums, const void* W, int ldw,
                                       const float* U, const float* S, const float* save_mean,
                                       const float* save_rstd, const float* gamma, float* dgamma,
                                       float* dbeta, void* dW, int lddw, void* Ws, void* Wt, float* nbias,
                                       int N, int K, int M, void* stream) {
  int rc = aux_shape_ok(N, K, "convbn_bwd_finalize");
  if (rc) return rc;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  return dispatch_jpl(K, [&](auto jpl) {
    constexpr int JPL = decltype(jpl)::value;
    convbn_bwd_finalize_kernel<JPL><<<N / kWPB, 32 * kWPB, K * sizeof(float), st>>>(
        P, ldp, gsums, reinterpret_cast<const __nv_bfloat16*>(W), ldw, U, S, save_mean, save_rstd, gamma,
        dgamma, dbeta, reinterpret_cast<__nv_bfloat16*>(dW), lddw, reinterpret_cast<__nv_bfloat16*>(Ws),
        reinterpret_cast<__nv_bfloat16*>(Wt), nbias, N, K, static_cast<float>(M));
    return check_cuda(cudaGetLastError(), "convbn_bwd_finalize launch");
  });
}
