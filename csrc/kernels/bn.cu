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
#include <mutex>

#include "optim.cuh"

namespace cb2 {

constexpr int kBnThreads = 256;
// Partial sums of different CTAs land in kBnCopies replicas of the [2C] accumulator
// (replica = blockIdx % kBnCopies): same-address L2 atomics serialise, and with ~600 CTAs on one
// replica they - not HBM - set the kernel time (measured: 60 us floor per launch).
constexpr int kBnCopies = 8;

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
    float* dst = out + static_cast<size_t>(blockIdx.x % kBnCopies) * 2 * C;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      atomicAdd(dst + c8 + j, a[j]);
      atomicAdd(dst + C + c8 + j, b[j]);
    }
  }
}

// Sum the replicas for this thread's 8 channels; one row-lane does it, the others read smem.
__device__ __forceinline__ void bn_gather_sums(const float* __restrict__ sums, int C, int c8,
                                               int rlane, float* s_a, float* s_b,
                                               float (&a)[8], float (&b)[8]) {
  if (rlane == 0) {
#pragma unroll
    for (int j = 0; j < 8; ++j) a[j] = b[j] = 0.f;
#pragma unroll
    for (int k = 0; k < kBnCopies; ++k) {   // unrolled: all replica loads in flight together
      const float* src = sums + static_cast<size_t>(k) * 2 * C;
      const float4 a0 = *reinterpret_cast<const float4*>(src + c8);
      const float4 a1 = *reinterpret_cast<const float4*>(src + c8 + 4);
      const float4 b0 = *reinterpret_cast<const float4*>(src + C + c8);
      const float4 b1 = *reinterpret_cast<const float4*>(src + C + c8 + 4);
      a[0] += a0.x; a[1] += a0.y; a[2] += a0.z; a[3] += a0.w;
      a[4] += a1.x; a[5] += a1.y; a[6] += a1.z; a[7] += a1.w;
      b[0] += b0.x; b[1] += b0.y; b[2] += b0.z; b[3] += b0.w;
      b[4] += b1.x; b[5] += b1.y; b[6] += b1.z; b[7] += b1.w;
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) { s_a[c8 + j] = a[j]; s_b[c8 + j] = b[j]; }
  }
  __syncthreads();
#pragma unroll
  for (int j = 0; j < 8; ++j) { a[j] = s_a[c8 + j]; b[j] = s_b[c8 + j]; }
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
  // four rows in flight per thread
  for (; row + 3 * stride < M; row += 4 * stride) {
    const uint4 w0 = ld_stream_v4(x + static_cast<size_t>(row) * C + c8);
    const uint4 w1 = ld_stream_v4(x + static_cast<size_t>(row + stride) * C + c8);
    const uint4 w2 = ld_stream_v4(x + static_cast<size_t>(row + 2 * stride) * C + c8);
    const uint4 w3 = ld_stream_v4(x + static_cast<size_t>(row + 3 * stride) * C + c8);
    const Vec8 a = unpack8_bf16(w0), b = unpack8_bf16(w1), c = unpack8_bf16(w2), d = unpack8_bf16(w3);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      s[j] += (a.v[j] + b.v[j]) + (c.v[j] + d.v[j]);
      q[j] = fmaf(a.v[j], a.v[j], fmaf(b.v[j], b.v[j], fmaf(c.v[j], c.v[j], fmaf(d.v[j], d.v[j], q[j]))));
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
                float* __restrict__ save_mean, float* __restrict__ save_invstd,
                const float* __restrict__ mean_bias, int M, int C, float eps, float momentum,
                int relu, int training) {
  __shared__ float s_a[2048], s_b[2048];
  const int tpr = C >> 3, rows = kBnThreads / tpr;
  const int rlane = threadIdx.x / tpr, c8 = (threadIdx.x % tpr) * 8;
  float sc[8], sh[8], tot[8], totsq[8];
  const float invM = 1.f / static_cast<float>(M);
  if (training) bn_gather_sums(sums, C, c8, rlane, s_a, s_b, tot, totsq);
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int c = c8 + j;
    float mean, var;
    if (training) {
      mean = tot[j] * invM;
      var = fmaxf(totsq[j] * invM - mean * mean, 0.f);
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
      run_mean[c] = (1.f - momentum) * run_mean[c] +
                    momentum * (mean + (mean_bias ? mean_bias[c] : 0.f));
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
                     const float* __restrict__ save_invstd, const float* __restrict__ gamma,
                     const float* __restrict__ beta, float* __restrict__ sums, int M, int C,
                     int relu) {
  extern __shared__ float smem[];
  const int tpr = C >> 3, rows = kBnThreads / tpr;
  const int rlane = threadIdx.x / tpr, c8 = (threadIdx.x % tpr) * 8;
  // relu without a saved output: the mask is recomputed from x (y > 0 <=> x*sc + sh > 0), which
  // saves reading y - one of three activation-sized reads - for every BN+ReLU without shortcut
  const bool mask_from_x = relu && y == nullptr;
  const bool mask_from_y = relu && y != nullptr;
  float mean[8], istd[8], db[8], dg[8], sc[8], sh[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    mean[j] = save_mean[c8 + j];
    istd[j] = save_invstd[c8 + j];
    sc[j] = gamma[c8 + j] * istd[j];
    sh[j] = beta[c8 + j] - mean[j] * sc[j];
    db[j] = dg[j] = 0.f;
  }
  const int stride = gridDim.x * rows;
  int row = blockIdx.x * rows + rlane;
  // two rows (up to six 16-byte loads) in flight per thread
  for (; row + stride < M; row += 2 * stride) {
    const size_t i0 = static_cast<size_t>(row) * C + c8;
    const size_t i1 = static_cast<size_t>(row + stride) * C + c8;
    const uint4 wg0 = ld_stream_v4(dy + i0), wg1 = ld_stream_v4(dy + i1);
    const uint4 wx0 = ld_stream_v4(x + i0), wx1 = ld_stream_v4(x + i1);
    uint4 wy0 = make_uint4(0, 0, 0, 0), wy1 = wy0;
    if (mask_from_y) { wy0 = ld_stream_v4(y + i0); wy1 = ld_stream_v4(y + i1); }
    Vec8 g0 = unpack8_bf16(wg0), g1 = unpack8_bf16(wg1);
    const Vec8 x0 = unpack8_bf16(wx0), x1 = unpack8_bf16(wx1);
    if (mask_from_y) {
      const Vec8 y0 = unpack8_bf16(wy0), y1 = unpack8_bf16(wy1);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        g0.v[j] = y0.v[j] > 0.f ? g0.v[j] : 0.f;
        g1.v[j] = y1.v[j] > 0.f ? g1.v[j] : 0.f;
      }
    } else if (mask_from_x) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        g0.v[j] = fmaf(x0.v[j], sc[j], sh[j]) > 0.f ? g0.v[j] : 0.f;
        g1.v[j] = fmaf(x1.v[j], sc[j], sh[j]) > 0.f ? g1.v[j] : 0.f;
      }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      db[j] += g0.v[j] + g1.v[j];
      dg[j] = fmaf(g0.v[j], (x0.v[j] - mean[j]) * istd[j],
                   fmaf(g1.v[j], (x1.v[j] - mean[j]) * istd[j], dg[j]));
    }
  }
  for (; row < M; row += stride) {
    const size_t idx = static_cast<size_t>(row) * C + c8;
    Vec8 g = unpack8_bf16(ld_stream_v4(dy + idx));
    const Vec8 xv = unpack8_bf16(ld_stream_v4(x + idx));
    if (mask_from_y) {
      const Vec8 yv = unpack8_bf16(ld_stream_v4(y + idx));
#pragma unroll
      for (int j = 0; j < 8; ++j) g.v[j] = yv.v[j] > 0.f ? g.v[j] : 0.f;
    } else if (mask_from_x) {
#pragma unroll
      for (int j = 0; j < 8; ++j) g.v[j] = fmaf(xv.v[j], sc[j], sh[j]) > 0.f ? g.v[j] : 0.f;
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
                    const float* __restrict__ beta, const float* __restrict__ sums,
                    float* __restrict__ dbeta_out,
                    float* __restrict__ dgamma_out, __nv_bfloat16* __restrict__ dx,
                    __nv_bfloat16* __restrict__ dres, int M, int C, int relu) {
  __shared__ float s_a[2048], s_b[2048];
  const int tpr = C >> 3, rows = kBnThreads / tpr;
  const int rlane = threadIdx.x / tpr, c8 = (threadIdx.x % tpr) * 8;
  const float invM = 1.f / static_cast<float>(M);
  const bool mask_from_x = relu && y == nullptr;
  const bool mask_from_y = relu && y != nullptr;
  float mean[8], istd[8], k0[8], k1[8], k2[8], db[8], dg[8], sh[8];
  bn_gather_sums(sums, C, c8, rlane, s_a, s_b, db, dg);
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int c = c8 + j;
    mean[j] = save_mean[c];
    istd[j] = save_invstd[c];
    k0[j] = gamma[c] * istd[j];
    sh[j] = beta[c] - mean[j] * k0[j];
    k1[j] = db[j] * invM;        // dbeta / M
    k2[j] = dg[j] * invM;        // dgamma / M
    if (blockIdx.x == 0 && rlane == 0) {
      dbeta_out[c] = db[j];
      dgamma_out[c] = dg[j];
    }
  }
  const int stride = gridDim.x * rows;
  for (int row = blockIdx.x * rows + rlane; row < M; row += stride) {
    const size_t idx = static_cast<size_t>(row) * C + c8;
    Vec8 g = unpack8_bf16(ld_stream_v4(dy + idx));
    const Vec8 xv = unpack8_bf16(ld_stream_v4(x + idx));
    if (mask_from_y) {
      const Vec8 yv = unpack8_bf16(ld_stream_v4(y + idx));
#pragma unroll
      for (int j = 0; j < 8; ++j) g.v[j] = yv.v[j] > 0.f ? g.v[j] : 0.f;
    } else if (mask_from_x) {
#pragma unroll
      for (int j = 0; j < 8; ++j) g.v[j] = fmaf(xv.v[j], k0[j], sh[j]) > 0.f ? g.v[j] : 0.f;
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
  const long long cap = static_cast<long long>(sms) * 4;
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

extern "C" int cb2_bn_scratch_floats(int C) { return kBnCopies * 2 * C; }

extern "C" int cb2_bn_forward(const void* x, const void* residual, void* y,
                              float* sums /*[cb2_bn_scratch_floats(C)], zeroed*/,
                              const float* gamma, const float* beta, float* run_mean, float* run_var,
                              float* save_mean, float* save_invstd, const float* mean_bias, int M,
                              int C, float eps, float momentum, int relu, int training,
                              void* stream) {
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
      save_invstd, mean_bias, M, C, eps, momentum, relu, training);
  return check_cuda(cudaGetLastError(), "bn_forward launch");
}

extern "C" int cb2_bn_backward(const void* dy, const void* y, const void* x, const float* save_mean,
                               const float* save_invstd, const float* gamma, const float* beta,
                               float* sums /*[cb2_bn_scratch_floats(C)], zeroed*/,
                               float* dbeta /*out [C]*/, float* dgamma /*out [C]*/, void* dx,
                               void* dres /*nullable*/, int M, int C, int relu, void* stream) {
  int rc = bn_check(M, C, "bn_backward");
  if (rc) return rc;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int grid = bn_grid(M, C);
  const int rows = kBnThreads / (C >> 3);
  const size_t smem = rows > 1 ? static_cast<size_t>(2) * rows * C * sizeof(float) : 0;
  bn_bwd_reduce_kernel<<<grid, kBnThreads, smem, st>>>(
      reinterpret_cast<const __nv_bfloat16*>(dy), reinterpret_cast<const __nv_bfloat16*>(y),
      reinterpret_cast<const __nv_bfloat16*>(x), save_mean, save_invstd, gamma, beta, sums, M, C,
      relu);
  bn_bwd_apply_kernel<<<grid, kBnThreads, 0, st>>>(
      reinterpret_cast<const __nv_bfloat16*>(dy), reinterpret_cast<const __nv_bfloat16*>(y),
      reinterpret_cast<const __nv_bfloat16*>(x), save_mean, save_invstd, gamma, beta, sums, dbeta,
      dgamma,
      reinterpret_cast<__nv_bfloat16*>(dx), reinterpret_cast<__nv_bfloat16*>(dres), M, C, relu);
  return check_cuda(cudaGetLastError(), "bn_backward launch");
}

// ---------------------------------------------------------------------------------------------
// Batch-norm forward for callers that cannot allocate device memory.
//
// cb2_bn_forward needs a zero-filled scratch buffer from its caller.  Some users of this C ABI
// have nowhere to get one: a plug-in loaded into a serving process that handed the whole HBM to
// its own pool at start-up (any further cudaMalloc fails), a host written in plain C or another
// language that links the library without torch's caching allocator, or a steady-state loop that
// must not pay the device-wide synchronisation cudaMalloc / cudaFree imply.  For them the
// scratch lives in the library image itself.
//
// The array below is per-device storage that every GPU which loads this library keeps until the
// process exits: kBnStaticScratchFloats * 4 bytes = 128 KiB, sized for the largest supported
// channel count (C = 2048).  Its name is only a handle in host code - the kernels need the device
// address, which cudaGetSymbolAddress returns for the CURRENT device, so it is looked up once per
// device and cached.
// ---------------------------------------------------------------------------------------------
constexpr int kBnStaticMaxC = 2048;
constexpr int kBnStaticScratchFloats = cb2::kBnCopies * 2 * kBnStaticMaxC;
static __device__ __align__(256) float g_bn_static_scratch[kBnStaticScratchFloats];

namespace cb2 {

__global__ void bn_static_scratch_clear_kernel(float* __restrict__ scratch, int n) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
    scratch[i] = 0.f;
}

constexpr int kBnStaticMaxDevices = 64;
static float* g_bn_static_ptr[kBnStaticMaxDevices] = {nullptr};
static std::mutex g_bn_static_mu;

// Device address of the static scratch on the current device (set-up on first use per device).
static int bn_static_scratch_ptr(float** out) {
  int dev = -1;
  CB2_CUDA_OK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kBnStaticMaxDevices) {
    set_last_error("bn_forward_static: device ordinal %d is outside the supported range [0, %d)",
                   dev, kBnStaticMaxDevices);
    return CB2_ERR_UNSUPPORTED;
  }
  std::lock_guard<std::mutex> lock(g_bn_static_mu);
  if (g_bn_static_ptr[dev] == nullptr) {
    void* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_bn_static_scratch);
    if (e != cudaSuccess || p == nullptr) {
      set_last_error("bn_forward_static: cudaGetSymbolAddress failed on device %d: %s", dev,
                     cudaGetErrorString(e));
      return CB2_ERR_CUDA;
    }
    g_bn_static_ptr[dev] = static_cast<float*>(p);
  }
  *out = g_bn_static_ptr[dev];
  return 0;
}

}  // namespace cb2

extern "C" int cb2_bn_static_scratch_max_channels(void) { return kBnStaticMaxC; }

extern "C" int cb2_bn_forward_static(const void* x, const void* residual, void* y,
                                     const float* gamma, const float* beta, float* run_mean,
                                     float* run_var, float* save_mean, float* save_invstd,
                                     const float* mean_bias, int M, int C, float eps,
                                     float momentum, int relu, int training, void* stream) {
  const long long need_ll = static_cast<long long>(kBnCopies) * 2 * C;
  const int need = static_cast<int>(need_ll);
  if (C > kBnStaticMaxC || need_ll > kBnStaticScratchFloats) {
    set_last_error("bn_forward_static: C=%d needs %d scratch floats but the built-in scratch holds "
                   "%d (C <= %d); use cb2_bn_forward with caller-provided scratch instead",
                   C, need, kBnStaticScratchFloats, kBnStaticMaxC);
    return CB2_ERR_UNSUPPORTED;
  }
  int rc = bn_check(M, C, "bn_forward_static");
  if (rc) return rc;
  float* scratch = nullptr;
  rc = bn_static_scratch_ptr(&scratch);
  if (rc) return rc;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int grid = bn_grid(M, C);
  const int rows = kBnThreads / (C >> 3);
  const size_t smem = rows > 1 ? static_cast<size_t>(2) * rows * C * sizeof(float) : 0;
  if (training) {
    // the array still holds the previous call's partial sums: the part this call accumulates into
    // has to start from zero
    bn_static_scratch_clear_kernel<<<(need + 255) / 256, 256, 0, st>>>(scratch, need);
    bn_stats_kernel<<<grid, kBnThreads, smem, st>>>(reinterpret_cast<const __nv_bfloat16*>(x),
                                                    scratch, M, C);
  }
  bn_apply_kernel<<<grid, kBnThreads, 0, st>>>(
      reinterpret_cast<const __nv_bfloat16*>(x), reinterpret_cast<const __nv_bfloat16*>(residual),
      reinterpret_cast<__nv_bfloat16*>(y), scratch, gamma, beta, run_mean, run_var, save_mean,
      save_invstd, mean_bias, M, C, eps, momentum, relu, training);
  return check_cuda(cudaGetLastError(), "bn_forward_static launch");
}
