// Fused NHWC batch normalisation for B200: BN (+ residual add) (+ ReLU) in one memory pass after
// a one-pass statistics reduction; backward in two passes (reduce, apply) with the ReLU mask and
// the residual gradient folded in.  bf16 activations, fp32 statistics / parameters.
//
// Layout: x is [M, C] row-major (M = N*H*W, channels_last), C = 8, 16, 32 or a multiple of 64, at
// most 2048.  Each thread owns 8 consecutive channels and moves them with one 16-byte access; a
// CTA of 256 threads owns a 64-channel slab of a group of rows (see BnMap).
//
// Reference workload: the BatchNormalization layers of ResNet-50 (K8 in SURVEY.md section 2.5;
// core/tests/examples/call_run_within_script_with_keras_fit.py:79-83) which TensorFlow runs as
// separate cuDNN batch-norm, add and ReLU kernels.
#include <stdlib.h>

#include <mutex>

#include "optim.cuh"

namespace cb2 {

constexpr int kBnThreads = 256;
// Partial sums of different CTAs land in kBnCopies replicas of the [2C] accumulator
// (replica = blockIdx % kBnCopies): same-address L2 atomics serialise, and with ~600 CTAs on one
// replica they - not HBM - set the kernel time (measured: 60 us floor per launch).
constexpr int kBnCopies = 8;

// Work split.  The [M, C] matrix is cut into channel slabs of W = min(C, 64) channels (128 bytes
// of every row) and row groups; CTA b owns slab b % slabs and row group b / slabs, so CTAs that run
// at the same time still sweep contiguous memory.  Inside a CTA a thread keeps 8 fixed channels
// (one 16-byte access per row) and W/8 threads share a row, i.e. a warp touches whole 128-byte
// lines.  Because a CTA only ever sees W channels, the end-of-kernel reduction costs 2*W atomics per
// CTA whatever C is (a CTA that spans all of C needs 2*C: 1.2 M atomics for C = 1024, which - not
// the 25 MB of data - set the time of the 14x14 and 7x7 layers).
struct BnMap {
  int tpr;     // threads per row = W / 8  (1, 2, 4 or 8)
  int rows;    // rows per CTA iteration = 256 / tpr
  int w;       // slab width in channels
  int cbase;   // first channel of the slab
  int c8;      // first of this thread's 8 channels
  int rlane;   // this thread's row inside an iteration
  int row0;    // first row of this CTA
  int stride;  // row step between iterations
  int rg;      // row group
};

__device__ __forceinline__ BnMap bn_map(int C) {
  BnMap m;
  m.w = C < 64 ? C : 64;
  m.tpr = m.w >> 3;
  m.rows = kBnThreads / m.tpr;
  const int slabs = C / m.w;
  const int slab = blockIdx.x % slabs;
  m.rg = blockIdx.x / slabs;
  const int nrg = gridDim.x / slabs;
  m.cbase = slab * m.w;
  m.c8 = m.cbase + (threadIdx.x % m.tpr) * 8;
  m.rlane = threadIdx.x / m.tpr;
  m.row0 = m.rg * m.rows + m.rlane;
  m.stride = nrg * m.rows;
  return m;
}

// CTA-wide sum of (a, b) over rows, then one atomic per channel and quantity into replica
// rg % kBnCopies of the [2C] accumulator.
__device__ __forceinline__ void bn_block_reduce_to_global(float (&a)[8], float (&b)[8],
                                                          const BnMap& m, int C,
                                                          float* __restrict__ out) {
  __shared__ float s_part[kBnThreads / 32][2][64];
  // lanes that differ in bits >= log2(tpr) hold other rows of the same channels
  for (int off = 16; off >= m.tpr; off >>= 1) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      a[j] += __shfl_xor_sync(0xffffffffu, a[j], off);
      b[j] += __shfl_xor_sync(0xffffffffu, b[j], off);
    }
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane < m.tpr) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      s_part[warp][0][lane * 8 + j] = a[j];
      s_part[warp][1][lane * 8 + j] = b[j];
    }
  }
  __syncthreads();
  if (threadIdx.x < 2 * m.w) {
    const int which = threadIdx.x / m.w, c = threadIdx.x % m.w;
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < kBnThreads / 32; ++w) t += s_part[w][which][c];
    atomicAdd(out + static_cast<size_t>(m.rg % kBnCopies) * 2 * C + which * C + m.cbase + c, t);
  }
}

// Totals of this thread's 8 channels: 2*W threads add up the replicas, the rest read smem.
__device__ __forceinline__ void bn_gather_sums(const float* __restrict__ sums, int C,
                                               const BnMap& m, float (&a)[8], float (&b)[8]) {
  __shared__ float s_tot[2][64];
  if (threadIdx.x < 2 * m.w) {
    const int which = threadIdx.x / m.w, c = threadIdx.x % m.w;
    float v[kBnCopies];
#pragma unroll
    for (int k = 0; k < kBnCopies; ++k)      // unrolled: all replica loads in flight together
      v[k] = sums[static_cast<size_t>(k) * 2 * C + which * C + m.cbase + c];
    float t = 0.f;
#pragma unroll
    for (int k = 0; k < kBnCopies; ++k) t += v[k];
    s_tot[which][c] = t;
  }
  __syncthreads();
  const int cl = (threadIdx.x % m.tpr) * 8;
#pragma unroll
  for (int j = 0; j < 8; ++j) { a[j] = s_tot[0][cl + j]; b[j] = s_tot[1][cl + j]; }
}

// sums[0:C] += sum_m x ; sums[C:2C] += sum_m x^2
__global__ void __launch_bounds__(kBnThreads)
bn_stats_kernel(const __nv_bfloat16* __restrict__ x, float* __restrict__ sums, int M, int C) {
  const BnMap m = bn_map(C);
  float s[8], q[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) s[j] = q[j] = 0.f;
  const int stride = m.stride;
  const __nv_bfloat16* xp = x + m.c8;
  int row = m.row0;
  // four rows in flight per thread
  for (; row + 3 * stride < M; row += 4 * stride) {
    const uint4 w0 = ld_stream_v4(xp + static_cast<size_t>(row) * C);
    const uint4 w1 = ld_stream_v4(xp + static_cast<size_t>(row + stride) * C);
    const uint4 w2 = ld_stream_v4(xp + static_cast<size_t>(row + 2 * stride) * C);
    const uint4 w3 = ld_stream_v4(xp + static_cast<size_t>(row + 3 * stride) * C);
    const Vec8 a = unpack8_bf16(w0), b = unpack8_bf16(w1), c = unpack8_bf16(w2), d = unpack8_bf16(w3);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      s[j] += (a.v[j] + b.v[j]) + (c.v[j] + d.v[j]);
      q[j] = fmaf(a.v[j], a.v[j], fmaf(b.v[j], b.v[j], fmaf(c.v[j], c.v[j], fmaf(d.v[j], d.v[j], q[j]))));
    }
  }
  for (; row < M; row += stride) {
    const Vec8 a = unpack8_bf16(ld_stream_v4(xp + static_cast<size_t>(row) * C));
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      s[j] += a.v[j];
      q[j] = fmaf(a.v[j], a.v[j], q[j]);
    }
  }
  bn_block_reduce_to_global(s, q, m, C, sums);
}

// g = (dy (+ dy2)) where the saved sign bit is set, else 0; sums[0:C] += sum_m g.  The masked
// gradient of a BN + shortcut + ReLU output in the form the folded 1x1-convolution backward
// (convbn_tcgen05.cu) takes it: as a GEMM operand, with its column sums already reduced.
__global__ void __launch_bounds__(kBnThreads)
bn_mask_grad_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ dy2,
                    const uint8_t* __restrict__ mask, __nv_bfloat16* __restrict__ g,
                    float* __restrict__ sums, int M, int C) {
  const BnMap m = bn_map(C);
  float s[8], q[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) s[j] = q[j] = 0.f;
  auto one = [&](size_t i, const uint4& wa, const uint4& wb, uint32_t bits) {
    Vec8 v = unpack8_bf16(wa);
    if (dy2) {
      const Vec8 w = unpack8_bf16(wb);
#pragma unroll
      for (int j = 0; j < 8; ++j) v.v[j] += w.v[j];
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      if (!((bits >> j) & 1u)) v.v[j] = 0.f;
    }
    const uint4 packed = pack8_bf16(v);
    const Vec8 r = unpack8_bf16(packed);          // sums of what the GEMMs will read
#pragma unroll
    for (int j = 0; j < 8; ++j) s[j] += r.v[j];
    st_stream_v4(g + i, packed);
  };
  const uint4 z4 = make_uint4(0, 0, 0, 0);
  int row = m.row0;
  // four rows (up to eight 16-byte loads) in flight per thread
  for (; row + 3 * m.stride < M; row += 4 * m.stride) {
    size_t i[4];
    uint4 wa[4], wb[4];
    uint32_t bits[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      i[u] = static_cast<size_t>(row + u * m.stride) * C + m.c8;
      wa[u] = ld_stream_v4(dy + i[u]);
      wb[u] = dy2 ? ld_stream_v4(dy2 + i[u]) : z4;
      bits[u] = mask ? mask[static_cast<size_t>(row + u * m.stride) * (C >> 3) + (m.c8 >> 3)] : 0xffu;
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) one(i[u], wa[u], wb[u], bits[u]);
  }
  for (; row < M; row += m.stride) {
    const size_t i = static_cast<size_t>(row) * C + m.c8;
    one(i, ld_stream_v4(dy + i), dy2 ? ld_stream_v4(dy2 + i) : z4,
        mask ? mask[static_cast<size_t>(row) * (C >> 3) + (m.c8 >> 3)] : 0xffu);
  }
  bn_block_reduce_to_global(s, q, m, C, sums);
}

// One bit per channel: output > 0.  Saved by the forward pass of a BN + shortcut + ReLU so that the
// backward kernels read one byte per 8 channels instead of the 16 bytes of the saved output.
__device__ __forceinline__ uint8_t bn_sign_bits(const Vec8& v) {
  uint32_t b = 0;
#pragma unroll
  for (int j = 0; j < 8; ++j) b |= (v.v[j] > 0.f ? 1u : 0u) << j;
  return static_cast<uint8_t>(b);
}

__device__ __forceinline__ Vec8 bn_affine(const Vec8& v, const float (&sc)[8], const float (&sh)[8]) {
  Vec8 o;
#pragma unroll
  for (int j = 0; j < 8; ++j) o.v[j] = fmaf(v.v[j], sc[j], sh[j]);
  return o;
}

// y = act(x * scale + shift (+ residual)); training: scale/shift from the batch sums (and row
// group 0 refreshes the running statistics and the saved mean / inv-std); inference: from the
// running statistics.
__global__ void __launch_bounds__(kBnThreads)
bn_apply_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ res,
                __nv_bfloat16* __restrict__ y, const float* __restrict__ sums,
                const float* __restrict__ gamma, const float* __restrict__ beta,
                float* __restrict__ run_mean, float* __restrict__ run_var,
                float* __restrict__ save_mean, float* __restrict__ save_invstd,
                const float* __restrict__ mean_bias, uint8_t* __restrict__ mask_out, int M, int C,
                float eps, float momentum, int relu, int training) {
  const BnMap m = bn_map(C);
  float sc[8], sh[8], tot[8], totsq[8];
  const float invM = 1.f / static_cast<float>(M);
  if (training) bn_gather_sums(sums, C, m, tot, totsq);
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int c = m.c8 + j;
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
    if (training && m.rg == 0 && m.rlane == 0) {
      save_mean[c] = mean;
      save_invstd[c] = istd;
      const float unbiased = M > 1 ? var * static_cast<float>(M) / static_cast<float>(M - 1) : var;
      run_mean[c] = (1.f - momentum) * run_mean[c] +
                    momentum * (mean + (mean_bias ? mean_bias[c] : 0.f));
      run_var[c] = (1.f - momentum) * run_var[c] + momentum * unbiased;
    }
  }
  const float lo = relu ? 0.f : -INFINITY;
  const int stride = m.stride;
  int row = m.row0;
  // two rows (up to four 16-byte loads) in flight per thread; y is re-read by the next conv, so
  // it is written with default caching (stays in L2)
  for (; row + stride < M; row += 2 * stride) {
    const size_t i0 = static_cast<size_t>(row) * C + m.c8;
    const size_t i1 = static_cast<size_t>(row + stride) * C + m.c8;
    const uint4 wx0 = ld_stream_v4(x + i0), wx1 = ld_stream_v4(x + i1);
    Vec8 v0 = bn_affine(unpack8_bf16(wx0), sc, sh), v1 = bn_affine(unpack8_bf16(wx1), sc, sh);
    if (res) {
      const Vec8 r0 = unpack8_bf16(ld_stream_v4(res + i0)), r1 = unpack8_bf16(ld_stream_v4(res + i1));
#pragma unroll
      for (int j = 0; j < 8; ++j) { v0.v[j] += r0.v[j]; v1.v[j] += r1.v[j]; }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) { v0.v[j] = fmaxf(v0.v[j], lo); v1.v[j] = fmaxf(v1.v[j], lo); }
    *reinterpret_cast<uint4*>(y + i0) = pack8_bf16(v0);
    *reinterpret_cast<uint4*>(y + i1) = pack8_bf16(v1);
    if (mask_out) {
      mask_out[static_cast<size_t>(row) * (C >> 3) + (m.c8 >> 3)] = bn_sign_bits(v0);
      mask_out[static_cast<size_t>(row + stride) * (C >> 3) + (m.c8 >> 3)] = bn_sign_bits(v1);
    }
  }
  for (; row < M; row += stride) {
    const size_t idx = static_cast<size_t>(row) * C + m.c8;
    Vec8 v = bn_affine(unpack8_bf16(ld_stream_v4(x + idx)), sc, sh);
    if (res) {
      const Vec8 r = unpack8_bf16(ld_stream_v4(res + idx));
#pragma unroll
      for (int j = 0; j < 8; ++j) v.v[j] += r.v[j];
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) v.v[j] = fmaxf(v.v[j], lo);
    *reinterpret_cast<uint4*>(y + idx) = pack8_bf16(v);
    if (mask_out) mask_out[static_cast<size_t>(row) * (C >> 3) + (m.c8 >> 3)] = bn_sign_bits(v);
  }
}

// g += second gradient stream (the output of this BN feeds two consumers: instead of an add
// kernel in front of the backward pass, both gradients are read here and summed in fp32)
__device__ __forceinline__ void bn_add8(Vec8& g, const uint4& w) {
  const Vec8 h = unpack8_bf16(w);
#pragma unroll
  for (int j = 0; j < 8; ++j) g.v[j] += h.v[j];
}

// dz = dy * (y > 0) when a ReLU follows the BN.  Without a saved output the mask is recomputed
// from x (y > 0 <=> x*sc + sh > 0), which saves reading y - one of three activation-sized reads -
// for every BN+ReLU without shortcut.
template <int MODE /*0 none, 1 from y, 2 from x, 3 from the sign bits saved by the forward pass*/>
__device__ __forceinline__ void bn_mask(Vec8& g, const Vec8& xv, const uint4& wy, const float (&sc)[8],
                                        const float (&sh)[8]) {
  if (MODE == 1) {
    const Vec8 yv = unpack8_bf16(wy);
#pragma unroll
    for (int j = 0; j < 8; ++j) g.v[j] = yv.v[j] > 0.f ? g.v[j] : 0.f;
  } else if (MODE == 2) {
#pragma unroll
    for (int j = 0; j < 8; ++j) g.v[j] = fmaf(xv.v[j], sc[j], sh[j]) > 0.f ? g.v[j] : 0.f;
  } else if (MODE == 3) {
#pragma unroll
    for (int j = 0; j < 8; ++j) g.v[j] = ((wy.x >> j) & 1u) ? g.v[j] : 0.f;
  }
}

// sums[0:C] += sum dz ; sums[C:2C] += sum dz * xhat
template <int MODE, bool DY2>
__global__ void __launch_bounds__(kBnThreads)
bn_bwd_reduce_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ dy2,
                     const __nv_bfloat16* __restrict__ y,
                     const __nv_bfloat16* __restrict__ x, const float* __restrict__ save_mean,
                     const float* __restrict__ save_invstd, const float* __restrict__ gamma,
                     const float* __restrict__ beta, float* __restrict__ sums, int M, int C) {
  const BnMap m = bn_map(C);
  // xhat = x * istd + nm  (nm = -mean * istd)
  float istd[8], nm[8], db[8], dg[8], sc[8], sh[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int c = m.c8 + j;
    istd[j] = save_invstd[c];
    nm[j] = -save_mean[c] * istd[j];
    sc[j] = gamma[c] * istd[j];
    sh[j] = fmaf(gamma[c], nm[j], beta[c]);
    db[j] = dg[j] = 0.f;
  }
  const int stride = m.stride;
  int row = m.row0;
  // two rows (up to six 16-byte loads) in flight per thread
  for (; row + stride < M; row += 2 * stride) {
    const size_t i0 = static_cast<size_t>(row) * C + m.c8;
    const size_t i1 = static_cast<size_t>(row + stride) * C + m.c8;
    const uint4 wg0 = ld_stream_v4(dy + i0), wg1 = ld_stream_v4(dy + i1);
    const uint4 wx0 = ld_stream_v4(x + i0), wx1 = ld_stream_v4(x + i1);
    uint4 wy0 = make_uint4(0, 0, 0, 0), wy1 = wy0, wh0 = wy0, wh1 = wy0;
    if (MODE == 1) { wy0 = ld_stream_v4(y + i0); wy1 = ld_stream_v4(y + i1); }
    if (MODE == 3) {            // `y` carries the saved sign bits: one byte per 8 channels
      const uint8_t* mk = reinterpret_cast<const uint8_t*>(y);
      wy0.x = mk[static_cast<size_t>(row) * (C >> 3) + (m.c8 >> 3)];
      wy1.x = mk[static_cast<size_t>(row + stride) * (C >> 3) + (m.c8 >> 3)];
    }
    if (DY2) { wh0 = ld_stream_v4(dy2 + i0); wh1 = ld_stream_v4(dy2 + i1); }
    Vec8 g0 = unpack8_bf16(wg0), g1 = unpack8_bf16(wg1);
    if (DY2) { bn_add8(g0, wh0); bn_add8(g1, wh1); }
    const Vec8 x0 = unpack8_bf16(wx0), x1 = unpack8_bf16(wx1);
    bn_mask<MODE>(g0, x0, wy0, sc, sh);
    bn_mask<MODE>(g1, x1, wy1, sc, sh);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      db[j] += g0.v[j] + g1.v[j];
      dg[j] = fmaf(g0.v[j], fmaf(x0.v[j], istd[j], nm[j]),
                   fmaf(g1.v[j], fmaf(x1.v[j], istd[j], nm[j]), dg[j]));
    }
  }
  for (; row < M; row += stride) {
    const size_t idx = static_cast<size_t>(row) * C + m.c8;
    Vec8 g = unpack8_bf16(ld_stream_v4(dy + idx));
    if (DY2) bn_add8(g, ld_stream_v4(dy2 + idx));
    const Vec8 xv = unpack8_bf16(ld_stream_v4(x + idx));
    uint4 wy = make_uint4(0, 0, 0, 0);
    if (MODE == 1) wy = ld_stream_v4(y + idx);
    if (MODE == 3) wy.x = reinterpret_cast<const uint8_t*>(y)[static_cast<size_t>(row) * (C >> 3) + (m.c8 >> 3)];
    bn_mask<MODE>(g, xv, wy, sc, sh);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      db[j] += g.v[j];
      dg[j] = fmaf(g.v[j], fmaf(xv.v[j], istd[j], nm[j]), dg[j]);
    }
  }
  bn_block_reduce_to_global(db, dg, m, C, sums);
}

// dx = gamma * istd * (dz - dbeta/M - xhat * dgamma/M) ; optionally dres = dz
template <int MODE, bool DY2>
__global__ void __launch_bounds__(kBnThreads)
bn_bwd_apply_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ dy2,
                    const __nv_bfloat16* __restrict__ y,
                    const __nv_bfloat16* __restrict__ x, const float* __restrict__ save_mean,
                    const float* __restrict__ save_invstd, const float* __restrict__ gamma,
                    const float* __restrict__ beta, const float* __restrict__ sums,
                    float* __restrict__ dbeta_out, float* __restrict__ dgamma_out,
                    __nv_bfloat16* __restrict__ dx, __nv_bfloat16* __restrict__ dres, int M, int C) {
  const BnMap m = bn_map(C);
  const float invM = 1.f / static_cast<float>(M);
  // dx = k0*dz - kx*x - kc   with kx = k0*k2*istd, kc = k0*(k1 + k2*nm)
  float k0[8], kx[8], kc[8], sc[8], sh[8], db[8], dg[8];
  bn_gather_sums(sums, C, m, db, dg);
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int c = m.c8 + j;
    const float istd = save_invstd[c];
    const float nm = -save_mean[c] * istd;
    const float g = gamma[c];
    k0[j] = g * istd;
    sc[j] = k0[j];
    sh[j] = fmaf(g, nm, beta[c]);
    const float k1 = db[j] * invM;        // dbeta / M
    const float k2 = dg[j] * invM;        // dgamma / M
    kx[j] = k0[j] * k2 * istd;
    kc[j] = k0[j] * fmaf(k2, nm, k1);
    if (m.rg == 0 && m.rlane == 0) {
      dbeta_out[c] = db[j];
      dgamma_out[c] = dg[j];
    }
  }
  const int stride = m.stride;
  int row = m.row0;
  for (; row + stride < M; row += 2 * stride) {
    const size_t i0 = static_cast<size_t>(row) * C + m.c8;
    const size_t i1 = static_cast<size_t>(row + stride) * C + m.c8;
    const uint4 wg0 = ld_stream_v4(dy + i0), wg1 = ld_stream_v4(dy + i1);
    const uint4 wx0 = ld_stream_v4(x + i0), wx1 = ld_stream_v4(x + i1);
    uint4 wy0 = make_uint4(0, 0, 0, 0), wy1 = wy0, wh0 = wy0, wh1 = wy0;
    if (MODE == 1) { wy0 = ld_stream_v4(y + i0); wy1 = ld_stream_v4(y + i1); }
    if (MODE == 3) {            // `y` carries the saved sign bits: one byte per 8 channels
      const uint8_t* mk = reinterpret_cast<const uint8_t*>(y);
      wy0.x = mk[static_cast<size_t>(row) * (C >> 3) + (m.c8 >> 3)];
      wy1.x = mk[static_cast<size_t>(row + stride) * (C >> 3) + (m.c8 >> 3)];
    }
    if (DY2) { wh0 = ld_stream_v4(dy2 + i0); wh1 = ld_stream_v4(dy2 + i1); }
    Vec8 g0 = unpack8_bf16(wg0), g1 = unpack8_bf16(wg1);
    if (DY2) { bn_add8(g0, wh0); bn_add8(g1, wh1); }
    const Vec8 x0 = unpack8_bf16(wx0), x1 = unpack8_bf16(wx1);
    bn_mask<MODE>(g0, x0, wy0, sc, sh);
    bn_mask<MODE>(g1, x1, wy1, sc, sh);
    if (dres) {
      *reinterpret_cast<uint4*>(dres + i0) = pack8_bf16(g0);
      *reinterpret_cast<uint4*>(dres + i1) = pack8_bf16(g1);
    }
    Vec8 o0, o1;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      o0.v[j] = fmaf(k0[j], g0.v[j], -fmaf(kx[j], x0.v[j], kc[j]));
      o1.v[j] = fmaf(k0[j], g1.v[j], -fmaf(kx[j], x1.v[j], kc[j]));
    }
    *reinterpret_cast<uint4*>(dx + i0) = pack8_bf16(o0);
    *reinterpret_cast<uint4*>(dx + i1) = pack8_bf16(o1);
  }
  for (; row < M; row += stride) {
    const size_t idx = static_cast<size_t>(row) * C + m.c8;
    Vec8 g = unpack8_bf16(ld_stream_v4(dy + idx));
    if (DY2) bn_add8(g, ld_stream_v4(dy2 + idx));
    const Vec8 xv = unpack8_bf16(ld_stream_v4(x + idx));
    uint4 wy = make_uint4(0, 0, 0, 0);
    if (MODE == 1) wy = ld_stream_v4(y + idx);
    if (MODE == 3) wy.x = reinterpret_cast<const uint8_t*>(y)[static_cast<size_t>(row) * (C >> 3) + (m.c8 >> 3)];
    bn_mask<MODE>(g, xv, wy, sc, sh);
    if (dres) *reinterpret_cast<uint4*>(dres + idx) = pack8_bf16(g);
    Vec8 o;
#pragma unroll
    for (int j = 0; j < 8; ++j) o.v[j] = fmaf(k0[j], g.v[j], -fmaf(kx[j], xv.v[j], kc[j]));
    *reinterpret_cast<uint4*>(dx + idx) = pack8_bf16(o);
  }
}

// Grid: a multiple of the slab count, about four CTAs per SM, never more row groups than there
// are CTA-iterations of rows.
static int bn_grid(int M, int C, bool backward = false) {
  const int w = C < 64 ? C : 64;
  const int slabs = C / w;
  const int rows = kBnThreads / (w >> 3);
  int sms = 148;
  int dev = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  long long nrg = (static_cast<long long>(M) + rows - 1) / rows;
  // CTAs per SM: measured on B200 over the ResNet-50 shapes (bench/bn_probe.py) - the one- and
  // two-stream forward kernels peak at 4 (6.0 TB/s on 411 MB), the backward kernels at 6 (6.1)
  static const int env_per_sm = [] {
    const char* e = getenv("TFC_BN_CTAS_PER_SM");
    return e ? atoi(e) : 0;
  }();
  const int per_sm = env_per_sm > 0 ? env_per_sm : (backward ? 6 : 4);
  long long cap = (static_cast<long long>(sms) * per_sm) / slabs;
  if (cap < 1) cap = 1;
  if (nrg > cap) nrg = cap;
  if (nrg < 1) nrg = 1;
  return static_cast<int>(nrg * slabs);
}

static bool bn_shape_ok(int C) {
  if (C <= 0 || (C & 7) || C > 2048) return false;
  return C % 64 == 0 || C == 8 || C == 16 || C == 32;
}

static int bn_check(int M, int C, const char* what) {
  if (M <= 0 || !bn_shape_ok(C)) {
    set_last_error("%s: unsupported shape M=%d C=%d (C must be 8, 16, 32 or a multiple of 64, <= 2048)",
                   what, M, C);
    return CB2_ERR_UNSUPPORTED;
  }
  return 0;
}

}  // namespace cb2

using namespace cb2;

extern "C" int cb2_bn_supported(int C) { return bn_shape_ok(C) ? 1 : 0; }

extern "C" int cb2_bn_scratch_floats(int C) { return kBnCopies * 2 * C; }

extern "C" int cb2_bn_stats(const void* x, float* sums, int M, int C, void* stream) {
  int rc = bn_check(M, C, "bn_stats");
  if (rc) return rc;
  bn_stats_kernel<<<bn_grid(M, C), kBnThreads, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const __nv_bfloat16*>(x), sums, M, C);
  return check_cuda(cudaGetLastError(), "bn_stats launch");
}

extern "C" int cb2_bn_mask_grad(const void* dy, const void* dy2, const void* mask, void* g, float* sums,
                                int M, int C, void* stream) {
  int rc = bn_check(M, C, "bn_mask_grad");
  if (rc) return rc;
  bn_mask_grad_kernel<<<bn_grid(M, C, true), kBnThreads, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const __nv_bfloat16*>(dy), reinterpret_cast<const __nv_bfloat16*>(dy2),
      reinterpret_cast<const uint8_t*>(mask), reinterpret_cast<__nv_bfloat16*>(g), sums, M, C);
  return check_cuda(cudaGetLastError(), "bn_mask_grad launch");
}

extern "C" int cb2_bn_forward_mask(const void* x, const void* residual, void* y, float* sums,
                                   const float* gamma, const float* beta, float* run_mean,
                                   float* run_var, float* save_mean, float* save_invstd,
                                   const float* mean_bias, void* mask_out, int M, int C, float eps,
                                   float momentum, int relu, int training, void* stream);

extern "C" int cb2_bn_forward(const void* x, const void* residual, void* y,
                              float* sums /*[cb2_bn_scratch_floats(C)], zeroed*/,
                              const float* gamma, const float* beta, float* run_mean, float* run_var,
                              float* save_mean, float* save_invstd, const float* mean_bias, int M,
                              int C, float eps, float momentum, int relu, int training,
                              void* stream) {
  return cb2_bn_forward_mask(x, residual, y, sums, gamma, beta, run_mean, run_var, save_mean,
                             save_invstd, mean_bias, nullptr, M, C, eps, momentum, relu, training, stream);
}

extern "C" int cb2_bn_forward_mask(const void* x, const void* residual, void* y, float* sums,
                                   const float* gamma, const float* beta, float* run_mean,
                                   float* run_var, float* save_mean, float* save_invstd,
                                   const float* mean_bias, void* mask_out /*nullable: [M*C/8] sign bits of y*/,
                                   int M, int C, float eps, float momentum, int relu, int training,
                                   void* stream) {
  int rc = bn_check(M, C, "bn_forward");
  if (rc) return rc;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int grid = bn_grid(M, C);
  if (training == 1)     // training == 2: `sums` already holds the statistics (GEMM epilogue)
    bn_stats_kernel<<<grid, kBnThreads, 0, st>>>(reinterpret_cast<const __nv_bfloat16*>(x), sums, M, C);
  bn_apply_kernel<<<grid, kBnThreads, 0, st>>>(
      reinterpret_cast<const __nv_bfloat16*>(x), reinterpret_cast<const __nv_bfloat16*>(residual),
      reinterpret_cast<__nv_bfloat16*>(y), sums, gamma, beta, run_mean, run_var, save_mean,
      save_invstd, mean_bias, reinterpret_cast<uint8_t*>(mask_out), M, C, eps, momentum, relu, training);
  return check_cuda(cudaGetLastError(), "bn_forward launch");
}

extern "C" int cb2_bn_backward2(const void* dy, const void* dy2, const void* y, const void* x,
                                const float* save_mean, const float* save_invstd, const float* gamma,
                                const float* beta, float* sums, float* dbeta, float* dgamma, void* dx,
                                void* dres, int M, int C, int relu, void* stream);

extern "C" int cb2_bn_backward(const void* dy, const void* y, const void* x, const float* save_mean,
                               const float* save_invstd, const float* gamma, const float* beta,
                               float* sums /*[cb2_bn_scratch_floats(C)], zeroed*/,
                               float* dbeta /*out [C]*/, float* dgamma /*out [C]*/, void* dx,
                               void* dres /*nullable*/, int M, int C, int relu, void* stream) {
  return cb2_bn_backward2(dy, nullptr, y, x, save_mean, save_invstd, gamma, beta, sums, dbeta, dgamma,
                          dx, dres, M, C, relu, stream);
}

extern "C" int cb2_bn_backward3(const void* dy, const void* dy2, const void* y, int y_is_mask,
                                const void* x, const float* save_mean, const float* save_invstd,
                                const float* gamma, const float* beta, float* sums, float* dbeta,
                                float* dgamma, void* dx, void* dres, int M, int C, int relu,
                                void* stream);

extern "C" int cb2_bn_backward2(const void* dy, const void* dy2 /*nullable: dy += dy2*/, const void* y,
                                const void* x, const float* save_mean, const float* save_invstd,
                                const float* gamma, const float* beta, float* sums, float* dbeta,
                                float* dgamma, void* dx, void* dres, int M, int C, int relu,
                                void* stream) {
  return cb2_bn_backward3(dy, dy2, y, 0, x, save_mean, save_invstd, gamma, beta, sums, dbeta, dgamma, dx,
                          dres, M, C, relu, stream);
}

extern "C" int cb2_bn_backward3(const void* dy, const void* dy2 /*nullable*/, const void* y /*nullable*/,
                                int y_is_mask /*y = the sign bits cb2_bn_forward_mask saved*/,
                                const void* x, const float* save_mean, const float* save_invstd,
                                const float* gamma, const float* beta, float* sums, float* dbeta,
                                float* dgamma, void* dx, void* dres, int M, int C, int relu,
                                void* stream) {
  int rc = bn_check(M, C, "bn_backward");
  if (rc) return rc;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int grid = bn_grid(M, C, true);
  const __nv_bfloat16* dyp = reinterpret_cast<const __nv_bfloat16*>(dy);
  const __nv_bfloat16* dy2p = reinterpret_cast<const __nv_bfloat16*>(dy2);
  const __nv_bfloat16* yp = reinterpret_cast<const __nv_bfloat16*>(y);
  const __nv_bfloat16* xp = reinterpret_cast<const __nv_bfloat16*>(x);
  __nv_bfloat16* dxp = reinterpret_cast<__nv_bfloat16*>(dx);
  __nv_bfloat16* drp = reinterpret_cast<__nv_bfloat16*>(dres);
  const int mode = !relu ? 0 : (y == nullptr ? 2 : (y_is_mask ? 3 : 1));
#define CB2_BN_BWD(MODE, DY2)                                                                      \
  bn_bwd_reduce_kernel<MODE, DY2><<<grid, kBnThreads, 0, st>>>(dyp, dy2p, yp, xp, save_mean,           \
                                                               save_invstd, gamma, beta, sums, M, C);  \
  bn_bwd_apply_kernel<MODE, DY2><<<grid, kBnThreads, 0, st>>>(dyp, dy2p, yp, xp, save_mean,            \
                                                              save_invstd, gamma, beta, sums, dbeta,   \
                                                              dgamma, dxp, drp, M, C);
  if (dy2p == nullptr) {
    if (mode == 0) { CB2_BN_BWD(0, false) }
    else if (mode == 1) { CB2_BN_BWD(1, false) }
    else if (mode == 2) { CB2_BN_BWD(2, false) }
    else { CB2_BN_BWD(3, false) }
  } else {
    if (mode == 0) { CB2_BN_BWD(0, true) }
    else if (mode == 1) { CB2_BN_BWD(1, true) }
    else if (mode == 2) { CB2_BN_BWD(2, true) }
    else { CB2_BN_BWD(3, true) }
  }
#undef CB2_BN_BWD
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
  if (training != 0 && training != 1) {
    // training == 2 of cb2_bn_forward means "the caller's scratch already holds the batch sums":
    // there is no caller scratch here
    set_last_error("bn_forward_static: training must be 0 (running statistics) or 1 (batch statistics), "
                   "got %d", training);
    return CB2_ERR_ARG;
  }
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  float* scratch = nullptr;
  if (training) {
    // The array is per device and its address is looked up for the CURRENT device: a stream that
    // belongs to another device would run the kernels there, on this device's address.
    int cur = -1, sdev = -1;
    CB2_CUDA_OK(cudaGetDevice(&cur));
    if (st != nullptr && st != cudaStreamLegacy && st != cudaStreamPerThread &&
        cudaStreamGetDevice(st, &sdev) == cudaSuccess && sdev != cur) {
      set_last_error("bn_forward_static: the stream belongs to device %d but device %d is current; make "
                     "the stream's device current before the call", sdev, cur);
      return CB2_ERR_ARG;
    }
    rc = bn_static_scratch_ptr(&scratch);      // inference reads no scratch: no lookup then
    if (rc) return rc;
  }
  const int grid = bn_grid(M, C);
  if (training) {
    // the array still holds the previous call's partial sums: the part this call accumulates into
    // has to start from zero
    bn_static_scratch_clear_kernel<<<(need + 255) / 256, 256, 0, st>>>(scratch, need);
    bn_stats_kernel<<<grid, kBnThreads, 0, st>>>(reinterpret_cast<const __nv_bfloat16*>(x),
                                                 scratch, M, C);
  }
  bn_apply_kernel<<<grid, kBnThreads, 0, st>>>(
      reinterpret_cast<const __nv_bfloat16*>(x), reinterpret_cast<const __nv_bfloat16*>(residual),
      reinterpret_cast<__nv_bfloat16*>(y), scratch, gamma, beta, run_mean, run_var, save_mean,
      save_invstd, mean_bias, nullptr, M, C, eps, momentum, relu, training);
  return check_cuda(cudaGetLastError(), "bn_forward_static launch");
}
