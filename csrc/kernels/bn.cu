// Fused NHWC batch normalisation for B200: BN (+ residual add) (+ ReLU) in one memory pass after
// a one-pass statistics reduction; backward in two passes (reduce, apply) with the ReLU mask and
// the residual gradient folded in.  bf16 activations, fp32 statistics / parameters.
//
// Layout: x is [M, C] row-major (M = N*H*W, channels_last), C a multiple of 8 and <= 2048.  A CTA
// of 256 threads covers 256/(C/8) rows per iteration; each thread owns 8 consecutive channels and
// moves them with one 16-byte access, so every warp reads/writes whole 128-byte lines.
//
// Reference workload: the BatchNormalization layers of ResNet-50 (K8 in SURVEY.md section 2.5;
// core/tests/examples/call_run_within_script_with_keras_fit.py:79-83) which TensorFlow runs as
// separate cuDNN batch-norm, add and ReLU kernels.
#include "optim.cuh"

namespace cb2 {

constexpr int kBnThreads = 256;

struct BnGeom {
  int tpr;       // threads per row = C / 8
  int rows;      // rows per CTA iteration = 256 / tpr
};

__device__ __forceinline__ void bn_block_reduce_to_global(float (&a)[8], float (&b)[8], int tpr,
                                                          int rows, int rlane, int c8, int C,
                                                          float* __restrict__ out, float* smem) {
  // smem: [rows][tpr*8] x 2
  if (rows > 1) {
    float* sa = smem;
    float* sb = smem + rows * tpr * 8;
    const int col = (threadIdx.x % tpr) * 8;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      sa[rlane * tpr * 8 + col + j] = a[j];
      sb[rlane * tpr * 8 + col + j] = b[j];
    }
    __syncthreads();
    if (rlane == 0) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float s = 0.f, q = 0.f;
        for (int r = 0; r < rows; ++r) {
          s += sa[r * tpr * 8 + col + j];
          q += sb[r * tpr * 8 + col + j];
        }
        a[j] = s;
        b[j] = q;
      }
    }
  }
  if (rlane == 0) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      atomicAdd(out + c8 + j, a[j]);
      atomicAdd(out + C + c8 + j, b[j]);
    }
  }
}

// sums[0:C] += sum_m x ; sums[C:2C] += sum_m x^2
__global__ void __launch_bounds__(kBnThreads)
bn_stats_kernel(const __nv_bfloat16* __restrict__ x, float* __restrict__ sums, int M, int C) {
  extern __shared__ float smem[];
  const int tpr = C >> 3, rows = kBnThreads / tpr;
  const int rlane = threadIdx.x / tpr, c8 = (threadIdx.x % tpr) * 8;
  float s[8], q[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) s[j] = q[j] = 0.f;
  const int stride = gridDim.x * rows;
  int row = blockIdx.x * rows + rlane;
  // two rows in flight per thread
  for (; row + stride < M; row += 2 * stride) {
    const uint4 w0 = ld_stream_v4(x + static_cast<size_t>(row) * C + c8);
    const uint4 w1 = ld_stream_v4(x + static_cast<size_t>(row + stride) * C + c8);
    const Vec8 a = unpack8_bf16(w0), b = unpack8_bf16(w1);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      s[j] += a.v[j] + b.v[j];
      q[j] = fmaf(a.v[j], a.v[j], fmaf(b.v[j], b.v[j], q[j]));
    }
  }
  for (; row < M; row += stride) {
    const Vec8 a = unpack8_bf16(ld_stream_v4(x + static_cast<size_t>(row) * C + c8));
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      s[j] += a.v[j];
      q[j] = fmaf(a.v[j], a.v[j], q[j]);
    }
  }
  bn_block_reduce_to_global(s, q, tpr, rows, rlane, c8, C, sums, smem);
}

// y = act(x * scale + shift (+ residual)); training: scale/shift from the batch sums (and block 0
// refreshes the running statistics and the saved mean / inv-std); inference: from running stats.
__global__ void __launch_bounds__(kBnThreads)
bn_apply_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ res,
                __nv_bfloat16* __restrict__ y, const float* __restrict__ sums,
                const float* __restrict__ gamma, const float* __restrict__ beta,
                float* __restrict__ run_mean, float* __restrict__ run_var,
                float* __restrict__ save_mean, float* __restrict__ save_invstd, int M, int C,
                float eps, float momentum, int relu, int training) {
  const int tpr = C >> 3, rows = kBnThreads / tpr;
  const int rlane = threadIdx.x / tpr, c8 = (threadIdx.x % tpr) * 8;
  float sc[8], sh[8];
  const float invM = 1.f / static_cast<float>(M);
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int c = c8 + j;
    float mean, var;
    if (training) {
      mean = sums[c] * invM;
      var = fmaxf(sums[C + c] * invM - mean * mean, 0.f);
    } else {
      mean = run_mean[c];
      var = run_var[c];
    }
    const float istd = rsqrtf(var + eps);
    sc[j] = gamma[c] * istd;
    sh[j] = beta[c] - mean * sc[j];
    if (training && blockIdx.x == 0 && rlane == 0) {
      save_mean[c] = mean;
      save_invstd[c] = istd;
      const float unbiased = M > 1 ? var * static_cast<float>(M) / static_cast<float>(M - 1) : var;
      run_mean[c] = (1.f - momentum) * run_mean[c] + momentum * mean;
      run_var[c] = (1.f - momentum) * run_var[c] + momentum * unbiased;
    }
  }
  const int stride = gridDim.x * rows;
  for (int row = blockIdx.x * rows + rlane; row < M; row += stride) {
    const size_t idx = static_cast<size_t>(row) * C + c8;
    Vec8 v = unpack8_bf16(ld_stream_v4(x + idx));
    if (res) {
      const Vec8 r = unpack8_bf16(ld_stream_v4(res + idx));
#pragma unroll
      for (int j = 0; j < 8; ++j) v.v[j] = fmaf(v.v[j], sc[j], sh[j]) + r.v[j];
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j) v.v[j] = fmaf(v.v[j], sc[j], sh[j]);
    }
    if (relu) {
#pragma unroll
      for (int j = 0; j < 8; ++j) v.v[j] = fmaxf(v.v[j], 0.f);
    }
    *reinterpret_cast<uint4*>(y + idx) = pack8_bf16(v);   // y is re-read by the next conv: keep in L2
  }
}

// sums[0:C] += sum dz ; sums[C:2C] += sum dz * xhat,  dz = dy * (y > 0 if relu)
__global__ void __launch_bounds__(kBnThreads)
bn_bwd_reduce_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ y,
                     const __nv_bfloat16* __restrict__ x, const float* __restrict__ save_mean,
                     const float* __restrict__ save_invstd, float* __restrict__ sums, int M, int C,
                     int relu) {
  extern __shared__ float smem[];
  const int tpr = C >> 3, rows = kBnThreads / tpr;
  const int rlane = threadIdx.x / tpr, c8 = (threadIdx.x % tpr) * 8;
  float mean[8], istd[8], db[8], dg[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    mean[j] = save_mean[c8 + j];
    istd[j] = save_invstd[c8 + j];
    db[j] = dg[j] = 0.f;
  }
  const int stride = gridDim.x * rows;
  for (int row = blockIdx.x * rows + rlane; row < M; row += stride) {
    const size_t idx = static_cast<size_t>(row) * C + c8;
    Vec8 g = unpack8_bf16(ld_stream_v4(dy + idx));
    const Vec8 xv = unpack8_bf16(ld_stream_v4(x + idx));
    if (relu) {
      const Vec8 yv = unpack8_bf16(ld_stream_v4(y + idx));
#pragma unroll
      for (int j = 0; j < 8; ++j) g.v[j] = yv.v[j] > 0.f ? g.v[j] : 0.f;
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      db[j] += g.v[j];
      dg[j] = fmaf(g.v[j], (xv.v[j] - mean[j]) * istd[j], dg[j]);
    }
  }
  bn_block_reduce_to_global(db, dg, tpr, rows, rlane, c8, C, sums, smem);
}

// dx = gamma * istd * (dz - dbeta/M - xhat * dgamma/M) ; optionally dres = dz
__global__ void __launch_bounds__(kBnThreads)
bn_bwd_apply_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ y,
                    const __nv_bfloat16* __restrict__ x, const float* __restrict__ save_mean,
                    const float* __restrict__ save_invstd, const float* __restrict__ gamma,
                    const float* __restrict__ sums, __nv_bfloat16* __restrict__ dx,
                    __nv_bfloat16* __restrict__ dres, int M, int C, int relu) {
  const int tpr = C >> 3, rows = kBnThreads / tpr;
  const int rlane = threadIdx.x / tpr, c8 = (threadIdx.x % tpr) * 8;
  const float invM = 1.f / static_cast<float>(M);
  float mean[8], istd[8], k0[8], k1[8], k2[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int c = c8 + j;
    mean[j] = save_mean[c];
    istd[j] = save_invstd[c];
    k0[j] = gamma[c] * istd[j];
    k1[j] = sums[c] * invM;        // dbeta / M
    k2[j] = sums[C + c] * invM;    // dgamma / M
  }
  const int stride = gridDim.x * rows;
  for (int row = blockIdx.x * rows + rlane; row < M; row += stride) {
    const size_t idx = static_cast<size_t>(row) * C + c8;
    Vec8 g = unpack8_bf16(ld_stream_v4(dy + idx));
    const Vec8 xv = unpack8_bf16(ld_stream_v4(x + idx));
    if (relu) {
      const Vec8 yv = unpack8_bf16(ld_stream_v4(y + idx));
#pragma unroll
      for (int j = 0; j < 8; ++j) g.v[j] = yv.v[j] > 0.f ? g.v[j] : 0.f;
    }
    if (dres) *reinterpret_cast<uint4*>(dres + idx) = pack8_bf16(g);
    Vec8 o;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float xh = (xv.v[j] - mean[j]) * istd[j];
      o.v[j] = k0[j] * (g.v[j] - k1[j] - xh * k2[j]);
    }
    *reinterpret_cast<uint4*>(dx + idx) = pack8_bf16(o);
  }
}

static int bn_grid(int M, int C) {
  const int rows = kBnThreads / (C >> 3);
  int sms = 148;
  int dev = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  long long want = (static_cast<long long>(M) + rows - 1) / rows;
  const long long cap = static_cast<long long>(sms) * 8;
  if (want > cap) want = cap;
  if (want < 1) want = 1;
  return static_cast<int>(want);
}

static int bn_check(int M, int C, const char* what) {
  if (M <= 0 || C <= 0 || (C & 7) || C > 2048 || (kBnThreads % (C >> 3)) != 0) {
    set_last_error("%s: unsupported shape M=%d C=%d (C must be a power-of-two multiple of 8, <= 2048)",
                   what, M, C);
    return CB2_ERR_UNSUPPORTED;
  }
  return 0;
}

}  // namespace cb2

using namespace cb2;

extern "C" int cb2_bn_supported(int C) {
  return (C > 0 && !(C & 7) && C <= 2048 && (kBnThreads % (C >> 3)) == 0) ? 1 : 0;
}

extern "C" int cb2_bn_forward(const void* x, const void* residual, void* y, float* sums /*[2C], zeroed*/,
                              const float* gamma, const float* beta, float* run_mean, float* run_var,
                              float* save_mean, float* save_invstd, int M, int C, float eps,
                              float momentum, int relu, int training, void* stream) {
  int rc = bn_check(M, C, "bn_forward");
  if (rc) return rc;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int grid = bn_grid(M, C);
  const int rows = kBnThreads / (C >> 3);
  const size_t smem = rows > 1 ? static_cast<size_t>(2) * rows * C * sizeof(float) : 0;
  if (training)
    bn_stats_kernel<<<grid, kBnThreads, smem, st>>>(reinterpret_cast<const __nv_bfloat16*>(x), sums, M, C);
  bn_apply_kernel<<<grid, kBnThreads, 0, st>>>(
      reinterpret_cast<const __nv_bfloat16*>(x), reinterpret_cast<const __nv_bfloat16*>(residual),
      reinterpret_cast<__nv_bfloat16*>(y), sums, gamma, beta, run_mean, run_var, save_mean,
      save_invstd, M, C, eps, momentum, relu, training);
  return check_cuda(cudaGetLastError(), "bn_forward launch");
}

extern "C" int cb2_bn_backward(const void* dy, const void* y, const void* x, const float* save_mean,
                               const float* save_invstd, const float* gamma,
                               float* sums /*[2C], zeroed: out dbeta | dgamma*/, void* dx,
                               void* dres /*nullable*/, int M, int C, int relu, void* stream) {
  int rc = bn_check(M, C, "bn_backward");
  if (rc) return rc;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int grid = bn_grid(M, C);
  const int rows = kBnThreads / (C >> 3);
  const size_t smem = rows > 1 ? static_cast<size_t>(2) * rows * C * sizeof(float) : 0;
  bn_bwd_reduce_kernel<<<grid, kBnThreads, smem, st>>>(
      reinterpret_cast<const __nv_bfloat16*>(dy), reinterpret_cast<const __nv_bfloat16*>(y),
      reinterpret_cast<const __nv_bfloat16*>(x), save_mean, save_invstd, sums, M, C, relu);
  bn_bwd_apply_kernel<<<grid, kBnThreads, 0, st>>>(
      reinterpret_cast<const __nv_bfloat16*>(dy), reinterpret_cast<const __nv_bfloat16*>(y),
      reinterpret_cast<const __nv_bfloat16*>(x), save_mean, save_invstd, gamma, sums,
      reinterpret_cast<__nv_bfloat16*>(dx), reinterpret_cast<__nv_bfloat16*>(dres), M, C, relu);
  return check_cuda(cudaGetLastError(), "bn_backward launch");
}
