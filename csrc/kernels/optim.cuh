// Optimizer update functors shared by the flat (single-GPU) optimizer kernel and the fused
// all-reduce epilogues.  One element at a time, all state in fp32.
#pragma once

#include "common.cuh"
#include "../../include/cloud_b200_kernels.h"

namespace cb2 {

struct Hyper {
  float lr, beta1, beta2, eps, wd, step, bc1, bc2, grad_scale, mom, nesterov, rms_mom;
  // derived
  float adam_keras_lr;  // lr * sqrt(bc2) / bc1
  float adam_torch_lr;  // lr / bc1
  float inv_sqrt_bc2;
};

__device__ __forceinline__ Hyper load_hyper(const float* __restrict__ h) {
  Hyper r;
  r.lr = h[0]; r.beta1 = h[1]; r.beta2 = h[2]; r.eps = h[3]; r.wd = h[4]; r.step = h[5];
  r.bc1 = h[6]; r.bc2 = h[7]; r.grad_scale = h[8]; r.mom = h[9]; r.nesterov = h[10];
  r.rms_mom = h[11];
  const float bc1 = r.bc1 > 0.f ? r.bc1 : 1.f;
  const float bc2 = r.bc2 > 0.f ? r.bc2 : 1.f;
  r.adam_keras_lr = r.lr * sqrtf(bc2) / bc1;
  r.adam_torch_lr = r.lr / bc1;
  r.inv_sqrt_bc2 = rsqrtf(bc2);
  return r;
}

// Returns the new parameter value; g is the already summed gradient (unscaled).
template <int OPT>
__device__ __forceinline__ float opt_update(float p, float g, float& s1, float& s2,
                                            const Hyper& h) {
  g *= h.grad_scale;
  if (OPT == CB2_OPT_NONE) {
    return g;
  } else if (OPT == CB2_OPT_ADAM_KERAS) {
    if (h.wd != 0.f) g = fmaf(h.wd, p, g);
    s1 = fmaf(h.beta1, s1, (1.f - h.beta1) * g);
    s2 = fmaf(h.beta2, s2, (1.f - h.beta2) * g * g);
    return p - h.adam_keras_lr * s1 / (sqrtf(s2) + h.eps);
  } else if (OPT == CB2_OPT_ADAM_TORCH || OPT == CB2_OPT_ADAMW) {
    if (OPT == CB2_OPT_ADAMW) p *= (1.f - h.lr * h.wd);
    else if (h.wd != 0.f) g = fmaf(h.wd, p, g);
    s1 = fmaf(h.beta1, s1, (1.f - h.beta1) * g);
    s2 = fmaf(h.beta2, s2, (1.f - h.beta2) * g * g);
    return p - h.adam_torch_lr * s1 / (sqrtf(s2) * h.inv_sqrt_bc2 + h.eps);
  } else if (OPT == CB2_OPT_SGD) {
    if (h.wd != 0.f) g = fmaf(h.wd, p, g);
    if (h.mom != 0.f) {
      s1 = fmaf(h.mom, s1, g);
      g = (h.nesterov != 0.f) ? fmaf(h.mom, s1, g) : s1;
    }
    return p - h.lr * g;
  } else {  // CB2_OPT_RMSPROP
    if (h.wd != 0.f) g = fmaf(h.wd, p, g);
    s1 = fmaf(h.mom, s1, (1.f - h.mom) * g * g);  // h.mom carries rho
    const float upd = h.lr * g * rsqrtf(s1 + h.eps);
    if (h.rms_mom != 0.f) {
      s2 = fmaf(h.rms_mom, s2, upd);
      return p - s2;
    }
    return p - upd;
  }
}

template <int OPT> struct OptTraits {
  static constexpr bool kS1 = OPT != CB2_OPT_NONE;
  static constexpr bool kS2 = OPT == CB2_OPT_ADAM_KERAS || OPT == CB2_OPT_ADAM_TORCH ||
                              OPT == CB2_OPT_ADAMW || OPT == CB2_OPT_RMSPROP;
};

// 8-element vectors ------------------------------------------------------------------------
struct Vec8 { float v[8]; };

__device__ __forceinline__ Vec8 load8_f32(const float* p) {
  Vec8 r;
  const float4 a = *reinterpret_cast<const float4*>(p);
  const float4 b = *reinterpret_cast<const float4*>(p + 4);
  r.v[0] = a.x; r.v[1] = a.y; r.v[2] = a.z; r.v[3] = a.w;
  r.v[4] = b.x; r.v[5] = b.y; r.v[6] = b.z; r.v[7] = b.w;
  return r;
}
__device__ __forceinline__ void store8_f32(float* p, const Vec8& r) {
  *reinterpret_cast<float4*>(p) = make_float4(r.v[0], r.v[1], r.v[2], r.v[3]);
  *reinterpret_cast<float4*>(p + 4) = make_float4(r.v[4], r.v[5], r.v[6], r.v[7]);
}
__device__ __forceinline__ Vec8 unpack8_bf16(const uint4& w) {
  Vec8 r;
  r.v[0] = bf16_lo(w.x); r.v[1] = bf16_hi(w.x); r.v[2] = bf16_lo(w.y); r.v[3] = bf16_hi(w.y);
  r.v[4] = bf16_lo(w.z); r.v[5] = bf16_hi(w.z); r.v[6] = bf16_lo(w.w); r.v[7] = bf16_hi(w.w);
  return r;
}
__device__ __forceinline__ uint4 pack8_bf16(const Vec8& r) {
  uint4 w;
  w.x = pack_bf16(r.v[0], r.v[1]); w.y = pack_bf16(r.v[2], r.v[3]);
  w.z = pack_bf16(r.v[4], r.v[5]); w.w = pack_bf16(r.v[6], r.v[7]);
  return w;
}

// Typed 8-element global access (T = float | __nv_bfloat16), element index i (multiple of 8).
template <typename T> __device__ __forceinline__ Vec8 load8(const T* base, size_t i);
template <> __device__ __forceinline__ Vec8 load8<float>(const float* base, size_t i) {
  Vec8 r;
  const uint4 a = ld_stream_v4(base + i);
  const uint4 b = ld_stream_v4(base + i + 4);
  r.v[0] = __uint_as_float(a.x); r.v[1] = __uint_as_float(a.y);
  r.v[2] = __uint_as_float(a.z); r.v[3] = __uint_as_float(a.w);
  r.v[4] = __uint_as_float(b.x); r.v[5] = __uint_as_float(b.y);
  r.v[6] = __uint_as_float(b.z); r.v[7] = __uint_as_float(b.w);
  return r;
}
template <> __device__ __forceinline__ Vec8 load8<__nv_bfloat16>(const __nv_bfloat16* base, size_t i) {
  return unpack8_bf16(ld_stream_v4(base + i));
}
template <typename T> __device__ __forceinline__ void store8(T* base, size_t i, const Vec8& r);
template <> __device__ __forceinline__ void store8<float>(float* base, size_t i, const Vec8& r) {
  uint4 a, b;
  a.x = __float_as_uint(r.v[0]); a.y = __float_as_uint(r.v[1]);
  a.z = __float_as_uint(r.v[2]); a.w = __float_as_uint(r.v[3]);
  b.x = __float_as_uint(r.v[4]); b.y = __float_as_uint(r.v[5]);
  b.z = __float_as_uint(r.v[6]); b.w = __float_as_uint(r.v[7]);
  st_stream_v4(base + i, a);
  st_stream_v4(base + i + 4, b);
}
template <> __device__ __forceinline__ void store8<__nv_bfloat16>(__nv_bfloat16* base, size_t i, const Vec8& r) {
  st_stream_v4(base + i, pack8_bf16(r));
}
template <typename T> __device__ __forceinline__ void zero8(T* base, size_t i) {
  const uint4 z = make_uint4(0, 0, 0, 0);
  st_stream_v4(base + i, z);
  if (sizeof(T) == 4) st_stream_v4(reinterpret_cast<float*>(base) + i + 4, z);
}

// Apply optimizer OPT to 8 consecutive elements.  `p_in` is the current parameter value (from the
// fp32 master if present, else from the parameter arena), gsum the reduced gradient.
template <int OPT>
__device__ __forceinline__ Vec8 update8(const Vec8& p_in, const Vec8& gsum, float* s1, float* s2,
                                        size_t si, const Hyper& h) {
  Vec8 a, b, out;
  if (OptTraits<OPT>::kS1) a = load8_f32(s1 + si);
  if (OptTraits<OPT>::kS2) b = load8_f32(s2 + si);
#pragma unroll
  for (int j = 0; j < 8; ++j) out.v[j] = opt_update<OPT>(p_in.v[j], gsum.v[j], a.v[j], b.v[j], h);
  if (OptTraits<OPT>::kS1) store8_f32(s1 + si, a);
  if (OptTraits<OPT>::kS2) store8_f32(s2 + si, b);
  return out;
}

}  // namespace cb2
