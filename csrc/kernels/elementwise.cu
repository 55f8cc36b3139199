// Memory-bound training kernels for B200: fused softmax cross-entropy (fwd+bwd+accuracy),
// ReLU-backward + bias-gradient, Philox dropout, uint8->bf16 input normalisation and the flat
// multi-optimizer update.  All are single-pass, 16-byte vectorised where the layout allows, with
// grids sized against the 148 SMs.
#include <algorithm>
#include <stdarg.h>
#include <stdio.h>

#include "optim.cuh"

namespace cb2 {

static thread_local char g_err[512] = "";

void set_last_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int check_cuda(cudaError_t e, const char* what) {
  if (e == cudaSuccess) return 0;
  set_last_error("CUDA error %d (%s) at %s", static_cast<int>(e), cudaGetErrorString(e), what);
  return CB2_ERR_CUDA;
}

static int sm_count() {
  static int n = 0;
  if (!n) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
  }
  return n;
}

// ---------------------------------------------------------------------------------------------
// Fused sparse softmax cross-entropy: one warp per row.
// ---------------------------------------------------------------------------------------------
template <typename T> __device__ __forceinline__ float to_f(T x);
template <> __device__ __forceinline__ float to_f<float>(float x) { return x; }
template <> __device__ __forceinline__ float to_f<__nv_bfloat16>(__nv_bfloat16 x) { return __bfloat162float(x); }
template <typename T> __device__ __forceinline__ T from_f(float x);
template <> __device__ __forceinline__ float from_f<float>(float x) { return x; }
template <> __device__ __forceinline__ __nv_bfloat16 from_f<__nv_bfloat16>(float x) { return __float2bfloat16(x); }

template <typename T>
__global__ void __launch_bounds__(256)
softmax_xent_kernel(const T* __restrict__ logits, const int64_t* __restrict__ labels,
                    T* __restrict__ dlogits, float* __restrict__ row_loss,
                    float* __restrict__ stats, int B, int C, int ld, float grad_scale,
                    float* __restrict__ loss_out, int add_count) {
  const int warps_per_block = blockDim.x >> 5;
  const int lane = threadIdx.x & 31;
  float loss_acc = 0.f, hit_acc = 0.f;
  for (int row = blockIdx.x * warps_per_block + (threadIdx.x >> 5); row < B;
       row += gridDim.x * warps_per_block) {
    const T* z = logits + static_cast<size_t>(row) * ld;
    const int label = static_cast<int>(labels[row]);
    // pass 1: max and arg-max (first maximal index wins, like argmax)
    float mx = -INFINITY;
    int amax = 0x7fffffff;
    for (int c = lane; c < C; c += 32) {
      const float v = to_f<T>(z[c]);
      if (v > mx) { mx = v; amax = c; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float om = __shfl_xor_sync(0xffffffffu, mx, o);
      const int oa = __shfl_xor_sync(0xffffffffu, amax, o);
      if (om > mx || (om == mx && oa < amax)) { mx = om; amax = oa; }
    }
    // pass 2: sum exp (row is L1/L2 resident by now)
    float s = 0.f;
    for (int c = lane; c < C; c += 32)
      s += exp2f((to_f<T>(z[c]) - mx) * 1.4426950408889634f);
    s = warp_sum(s);
    const float inv = 1.f / s;
    const float lse = mx + logf(s);
    // pass 3: gradient
    T* dz = dlogits + static_cast<size_t>(row) * ld;
    float zl = 0.f;
    for (int c = lane; c < C; c += 32) {
      const float v = to_f<T>(z[c]);
      float pr = exp2f((v - mx) * 1.4426950408889634f) * inv;
      if (c == label) { zl = v; pr -= 1.f; }
      dz[c] = from_f<T>(pr * grad_scale);
    }
    zl = warp_sum(zl);
    if (lane == 0) {
      const float l = lse - zl;
      if (row_loss) row_loss[row] = l;
      loss_acc += l;
      hit_acc += (amax == label) ? 1.f : 0.f;
    }
  }
  // one atomic pair per block
  __shared__ float sh[2][8];
  if (lane == 0) { sh[0][threadIdx.x >> 5] = loss_acc; sh[1][threadIdx.x >> 5] = hit_acc; }
  __syncthreads();
  if (threadIdx.x == 0) {
    float l = 0.f, h = 0.f;
    for (int w = 0; w < warps_per_block; ++w) { l += sh[0][w]; h += sh[1][w]; }
    if (stats) {
      atomicAdd(stats + 0, l);
      atomicAdd(stats + 1, h);
      if (add_count && blockIdx.x == 0) atomicAdd(stats + 2, static_cast<float>(B));
    }
    if (loss_out) atomicAdd(loss_out, l * grad_scale);     // this call's mean loss (zero on entry)
  }
}

// ---------------------------------------------------------------------------------------------
// dz = dy * (y > 0), db[n] (+)= sum_m dz[m, n].   Block = 32 columns x 8 row-lanes.
// ---------------------------------------------------------------------------------------------
template <typename DB>
__global__ void __launch_bounds__(256)
relu_bwd_bias_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ y,
                     __nv_bfloat16* __restrict__ dz, DB* __restrict__ db, int M, int N, int ld,
                     int row_chunk, int accumulate, float scale) {
  const int col = blockIdx.x * 32 + (threadIdx.x & 31);
  const int r_lane = threadIdx.x >> 5;
  const int r0 = blockIdx.y * row_chunk;
  const int r1 = min(M, r0 + row_chunk);
  float acc = 0.f;
  if (col < N) {
    for (int r = r0 + r_lane; r < r1; r += 8) {
      const size_t idx = static_cast<size_t>(r) * ld + col;
      float g = __bfloat162float(dy[idx]) * scale;
      if (y && !(__bfloat162float(y[idx]) > 0.f)) g = 0.f;
      if (dz) dz[idx] = __float2bfloat16(g);
      acc += g;
    }
  }
  __shared__ float sh[8][33];
  sh[r_lane][threadIdx.x & 31] = acc;
  __syncthreads();
  if (r_lane == 0 && col < N && db) {
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += sh[i][threadIdx.x & 31];
    if (gridDim.y == 1 && !accumulate) {
      db[col] = from_f<DB>(s);
    } else {
      // multi-chunk or accumulate: fp32 atomics (bf16 bias grads take the single-chunk path)
      atomicAdd(reinterpret_cast<float*>(db) + col, s);
    }
  }
}

__global__ void zero_f32_kernel(float* p, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = 0.f;
}

// ---------------------------------------------------------------------------------------------
// Philox-4x32-10 dropout; the mask is a pure function of (seed, offset + element index).
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint4 philox4x32(uint64_t ctr, uint64_t key, uint32_t stream_lo,
                                            uint32_t stream_hi) {
  uint32_t c0 = static_cast<uint32_t>(ctr), c1 = static_cast<uint32_t>(ctr >> 32);
  uint32_t c2 = stream_lo, c3 = stream_hi;
  uint32_t k0 = static_cast<uint32_t>(key), k1 = static_cast<uint32_t>(key >> 32);
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
    const uint32_t n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  return make_uint4(c0, c1, c2, c3);
}

template <typename T>
__global__ void __launch_bounds__(256)
dropout_kernel(const T* __restrict__ x, T* __restrict__ y, size_t n, float p, float scale,
               uint64_t seed, uint64_t offset, const float* __restrict__ step_ptr) {
  const uint32_t thresh = static_cast<uint32_t>(fminf(p, 1.f) * 4294967295.0f);
  // the training step lives in device memory so a CUDA-graph replay draws a fresh mask
  const uint32_t step = step_ptr ? static_cast<uint32_t>(*step_ptr) : 0u;
  for (size_t g = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; g * 4 < n;
       g += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const uint4 r = philox4x32(offset + g, seed, step, 0x5eedu);
    const uint32_t rr[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const size_t i = g * 4 + j;
      if (i < n) y[i] = from_f<T>(rr[j] >= thresh ? to_f<T>(x[i]) * scale : 0.f);
    }
  }
}

// ---------------------------------------------------------------------------------------------
// uint8 NHWC -> bf16 NHWC normalised.  One thread = 4 pixels of C=3 (12 bytes in, 24 out) for the
// common RGB case; generic path otherwise.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
u8_to_bf16_norm_kernel(const uint8_t* __restrict__ in, __nv_bfloat16* __restrict__ out,
                       size_t n_elems, int channels, float m0, float m1, float m2, float s0,
                       float s1, float s2) {
  const float mean[3] = {m0, m1, m2};
  const float istd[3] = {s0, s1, s2};
  if (channels == 3 && (n_elems % 12) == 0) {
    const size_t n_groups = n_elems / 12;
    for (size_t g = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; g < n_groups;
         g += static_cast<size_t>(gridDim.x) * blockDim.x) {
      const uint32_t* src = reinterpret_cast<const uint32_t*>(in + g * 12);
      const uint32_t w[3] = {__ldg(src), __ldg(src + 1), __ldg(src + 2)};
      float f[12];
#pragma unroll
      for (int j = 0; j < 12; ++j) {
        const float v = static_cast<float>((w[j >> 2] >> ((j & 3) * 8)) & 0xffu) * (1.f / 255.f);
        f[j] = (v - mean[j % 3]) * istd[j % 3];
      }
      uint32_t* dst = reinterpret_cast<uint32_t*>(out + g * 12);
#pragma unroll
      for (int j = 0; j < 6; ++j) dst[j] = pack_bf16(f[2 * j], f[2 * j + 1]);
    }
  } else {
    for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n_elems;
         i += static_cast<size_t>(gridDim.x) * blockDim.x) {
      const int c = static_cast<int>(i % channels) % 3;
      out[i] = __float2bfloat16((static_cast<float>(in[i]) * (1.f / 255.f) - mean[c]) * istd[c]);
    }
  }
}

// uint8 NHWC RGB image -> normalised bf16 in the space-to-depth layout of the ResNet stem
// (cb2_conv_stem_forward): out[b, r, s, (a*2+bb)*3 + c] = norm(in[b, 2r-4+a, 2s-4+bb, c]), zero
// outside the image (the convolution pads the NORMALISED image with zeros) and in channels
// 12..15.  One thread per output pixel: 2 x 6 input bytes in, 32 bytes out.
__global__ void __launch_bounds__(256)
u8_to_s2d_bf16_kernel(const uint8_t* __restrict__ in, __nv_bfloat16* __restrict__ out, int B, int H,
                      int W, float m0, float m1, float m2, float s0, float s1, float s2) {
  const int R = H / 2 + 3, S = W / 2 + 3;
  const unsigned total = static_cast<unsigned>(B) * R * S;
  const float mean[3] = {m0, m1, m2};
  const float istd[3] = {s0, s1, s2};
  for (unsigned t = blockIdx.x * blockDim.x + threadIdx.x; t < total; t += gridDim.x * blockDim.x) {
    const int s = static_cast<int>(t % S);
    unsigned q = t / S;
    const int r = static_cast<int>(q % R);
    const int b = static_cast<int>(q / R);
    float f[12];
#pragma unroll
    for (int j = 0; j < 12; ++j) f[j] = 0.f;
    const int w0 = 2 * s - 4;                       // even; both pixels of the pair in or out together
    if (w0 >= 0 && w0 < W) {
#pragma unroll
      for (int a = 0; a < 2; ++a) {
        const int h = 2 * r - 4 + a;
        if (h >= 0 && h < H) {
          const uint16_t* src = reinterpret_cast<const uint16_t*>(
              in + (static_cast<size_t>(b) * H * W + static_cast<size_t>(h) * W + w0) * 3);
          const uint32_t u0 = __ldg(src), u1 = __ldg(src + 1), u2 = __ldg(src + 2);
          const uint32_t bytes[6] = {u0 & 0xffu, u0 >> 8, u1 & 0xffu, u1 >> 8, u2 & 0xffu, u2 >> 8};
#pragma unroll
          for (int k = 0; k < 6; ++k)
            f[a * 6 + k] = (static_cast<float>(bytes[k]) * (1.f / 255.f) - mean[k % 3]) * istd[k % 3];
        }
      }
    }
    uint4* dst = reinterpret_cast<uint4*>(out + static_cast<size_t>(t) * 16);
    dst[0] = make_uint4(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]), pack_bf16(f[4], f[5]),
                        pack_bf16(f[6], f[7]));
    dst[1] = make_uint4(pack_bf16(f[8], f[9]), pack_bf16(f[10], f[11]), 0u, 0u);
  }
}

// ---------------------------------------------------------------------------------------------
// Optimizers
// ---------------------------------------------------------------------------------------------
__global__ void optim_tick_kernel(float* hyper) {
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    const float t = hyper[5] + 1.f;
    hyper[5] = t;
    hyper[6] = 1.f - powf(hyper[1], t);
    hyper[7] = 1.f - powf(hyper[2], t);
  }
}

template <int OPT, typename GT, typename PT>
__global__ void __launch_bounds__(512)
optim_flat_kernel(float* __restrict__ master, GT* __restrict__ grad, float* __restrict__ s1,
                  float* __restrict__ s2, PT* __restrict__ param_out, size_t n,
                  const float* __restrict__ hyper, int zero_grad) {
  const Hyper h = load_hyper(hyper);
  const size_t nvec = n / 8;
  for (size_t vi = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; vi < nvec;
       vi += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const size_t i = vi * 8;
    const Vec8 g = load8<GT>(grad, i);
    Vec8 p;
    if (master) p = load8_f32(master + i);
    else p = load8<PT>(param_out, i);
    const Vec8 pn = update8<OPT>(p, g, s1, s2, i, h);
    if (master) store8_f32(master + i, pn);
    if (param_out) store8<PT>(param_out, i, pn);
    if (zero_grad) zero8<GT>(grad, i);
  }
  // scalar tail
  const size_t tail0 = nvec * 8;
  for (size_t i = tail0 + blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const float g = to_f<GT>(grad[i]);
    float p = master ? master[i] : to_f<PT>(param_out[i]);
    float a = OptTraits<OPT>::kS1 ? s1[i] : 0.f, b = OptTraits<OPT>::kS2 ? s2[i] : 0.f;
    p = opt_update<OPT>(p, g, a, b, h);
    if (OptTraits<OPT>::kS1) s1[i] = a;
    if (OptTraits<OPT>::kS2) s2[i] = b;
    if (master) master[i] = p;
    if (param_out) param_out[i] = from_f<PT>(p);
    if (zero_grad) grad[i] = from_f<GT>(0.f);
  }
}

template <int OPT>
static int launch_optim_flat(float* master, void* grad, int gdt, float* s1, float* s2, void* pout,
                             int pdt, size_t n, const float* hyper, int zero_grad,
                             cudaStream_t st) {
  const size_t nvec = (n + 7) / 8;
  int blocks = static_cast<int>((nvec + 511) / 512);
  const int cap = sm_count() * 4;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
#define CB2_LAUNCH(GT, PT)                                                                  \
  optim_flat_kernel<OPT, GT, PT><<<blocks, 512, 0, st>>>(master, reinterpret_cast<GT*>(grad), \
                                                        s1, s2, reinterpret_cast<PT*>(pout), n, \
                                                        hyper, zero_grad)
  if (gdt == CB2_DTYPE_F32 && pdt == CB2_DTYPE_F32) CB2_LAUNCH(float, float);
  else if (gdt == CB2_DTYPE_BF16 && pdt == CB2_DTYPE_BF16) CB2_LAUNCH(__nv_bfloat16, __nv_bfloat16);
  else if (gdt == CB2_DTYPE_F32 && pdt == CB2_DTYPE_BF16) CB2_LAUNCH(float, __nv_bfloat16);
  else CB2_LAUNCH(__nv_bfloat16, float);
#undef CB2_LAUNCH
  return check_cuda(cudaGetLastError(), "optim_flat launch");
}

}  // namespace cb2

using namespace cb2;

extern "C" const char* cb2_last_error(void) { return g_err; }
extern "C" int cb2_version(void) { return 100; }

extern "C" int cb2_softmax_xent_ex(const void* logits, int dtype, const int64_t* labels, void* dlogits,
                                   float* row_loss, float* stats, int B, int C, int ld, float grad_scale,
                                   float* loss_out, int add_count, void* stream) {
  if (B <= 0 || C <= 0 || ld < C) {
    set_last_error("softmax_xent: bad shape B=%d C=%d ld=%d", B, C, ld);
    return CB2_ERR_ARG;
  }
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  int blocks = (B + 7) / 8;
  const int cap = sm_count() * 8;
  if (blocks > cap) blocks = cap;
  if (dtype == CB2_DTYPE_F32)
    softmax_xent_kernel<float><<<blocks, 256, 0, st>>>(
        reinterpret_cast<const float*>(logits), labels, reinterpret_cast<float*>(dlogits), row_loss,
        stats, B, C, ld, grad_scale, loss_out, add_count);
  else if (dtype == CB2_DTYPE_BF16)
    softmax_xent_kernel<__nv_bfloat16><<<blocks, 256, 0, st>>>(
        reinterpret_cast<const __nv_bfloat16*>(logits), labels,
        reinterpret_cast<__nv_bfloat16*>(dlogits), row_loss, stats, B, C, ld, grad_scale, loss_out, add_count);
  else {
    set_last_error("softmax_xent: unsupported dtype %d", dtype);
    return CB2_ERR_ARG;
  }
  return check_cuda(cudaGetLastError(), "softmax_xent launch");
}

extern "C" int cb2_softmax_xent(const void* logits, int dtype, const int64_t* labels, void* dlogits,
                                float* row_loss, float* stats, int B, int C, int ld, float grad_scale,
                                void* stream) {
  return cb2_softmax_xent_ex(logits, dtype, labels, dlogits, row_loss, stats, B, C, ld, grad_scale, nullptr,
                             0, stream);
}

extern "C" int cb2_relu_bwd_bias_grad_ex(const void* dy, const void* y, void* dz, void* db, int db_dtype,
                                         int M, int N, int ld, int accumulate, float scale, void* stream);

extern "C" int cb2_relu_bwd_bias_grad(const void* dy, const void* y, void* dz, void* db,
                                      int db_dtype, int M, int N, int ld, int accumulate,
                                      void* stream) {
  return cb2_relu_bwd_bias_grad_ex(dy, y, dz, db, db_dtype, M, N, ld, accumulate, 1.f, stream);
}

extern "C" int cb2_relu_bwd_bias_grad_ex(const void* dy, const void* y, void* dz, void* db, int db_dtype,
                                         int M, int N, int ld, int accumulate, float scale, void* stream) {
  if (M <= 0 || N <= 0 || ld < N) {
    set_last_error("relu_bwd_bias_grad: bad shape M=%d N=%d ld=%d", M, N, ld);
    return CB2_ERR_ARG;
  }
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int col_blocks = (N + 31) / 32;
  int row_chunks = 1;
  if (db_dtype == CB2_DTYPE_F32) {
    // enough CTAs to fill the chip, each covering >= 64 rows
    row_chunks = (sm_count() * 2 + col_blocks - 1) / col_blocks;
    const int max_chunks = (M + 63) / 64;
    if (row_chunks > max_chunks) row_chunks = max_chunks;
    if (row_chunks < 1) row_chunks = 1;
  } else if (accumulate) {
    set_last_error("relu_bwd_bias_grad: accumulate needs an fp32 db");
    return CB2_ERR_ARG;
  }
  const int row_chunk = (M + row_chunks - 1) / row_chunks;
  if (db && db_dtype == CB2_DTYPE_F32 && row_chunks > 1 && !accumulate)
    zero_f32_kernel<<<(N + 255) / 256, 256, 0, st>>>(reinterpret_cast<float*>(db), N);
  dim3 grid(col_blocks, row_chunks);
  if (db_dtype == CB2_DTYPE_F32)
    relu_bwd_bias_kernel<float><<<grid, 256, 0, st>>>(
        reinterpret_cast<const __nv_bfloat16*>(dy), reinterpret_cast<const __nv_bfloat16*>(y),
        reinterpret_cast<__nv_bfloat16*>(dz), reinterpret_cast<float*>(db), M, N, ld, row_chunk,
        accumulate, scale);
  else
    relu_bwd_bias_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>(
        reinterpret_cast<const __nv_bfloat16*>(dy), reinterpret_cast<const __nv_bfloat16*>(y),
        reinterpret_cast<__nv_bfloat16*>(dz), reinterpret_cast<__nv_bfloat16*>(db), M, N, ld,
        row_chunk, 0, scale);
  return check_cuda(cudaGetLastError(), "relu_bwd_bias_grad launch");
}

extern "C" int cb2_dropout(const void* x, void* y, size_t n, float p, uint64_t seed,
                           uint64_t offset, const float* step_ptr, int dtype, void* stream) {
  if (p < 0.f || p >= 1.f) {
    set_last_error("dropout: rate must be in [0,1), got %f", p);
    return CB2_ERR_ARG;
  }
  if (n == 0) return 0;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const size_t groups = (n + 3) / 4;
  int blocks = static_cast<int>((groups + 255) / 256);
  const int cap = sm_count() * 8;
  if (blocks > cap) blocks = cap;
  const float scale = 1.f / (1.f - p);
  if (dtype == CB2_DTYPE_F32)
    dropout_kernel<float><<<blocks, 256, 0, st>>>(reinterpret_cast<const float*>(x),
                                                  reinterpret_cast<float*>(y), n, p, scale, seed,
                                                  offset, step_ptr);
  else
    dropout_kernel<__nv_bfloat16><<<blocks, 256, 0, st>>>(
        reinterpret_cast<const __nv_bfloat16*>(x), reinterpret_cast<__nv_bfloat16*>(y), n, p, scale,
        seed, offset, step_ptr);
  return check_cuda(cudaGetLastError(), "dropout launch");
}

extern "C" int cb2_u8_to_bf16_normalize(const uint8_t* in, void* out, size_t n_pixels, int channels,
                                        const float* mean3, const float* inv_std3, void* stream) {
  if (channels <= 0) {
    set_last_error("u8_to_bf16_normalize: channels must be positive");
    return CB2_ERR_ARG;
  }
  const size_t n = n_pixels * channels;
  if (n == 0) return 0;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const size_t work = (channels == 3 && n % 12 == 0) ? n / 12 : n;
  int blocks = static_cast<int>((work + 255) / 256);
  const int cap = sm_count() * 8;
  if (blocks > cap) blocks = cap;
  u8_to_bf16_norm_kernel<<<blocks, 256, 0, st>>>(in, reinterpret_cast<__nv_bfloat16*>(out), n,
                                                 channels, mean3[0], mean3[1], mean3[2],
                                                 inv_std3[0], inv_std3[1], inv_std3[2]);
  return check_cuda(cudaGetLastError(), "u8_to_bf16_normalize launch");
}

extern "C" int cb2_u8_to_s2d_bf16(const uint8_t* in, void* out, int B, int H, int W,
                                  const float* mean3, const float* inv_std3, void* stream) {
  if (B <= 0 || H <= 0 || W <= 0 || (H & 1) || (W & 1) ||
      static_cast<long long>(B) * (H / 2 + 3) * (W / 2 + 3) >= (1LL << 31)) {
    set_last_error("u8_to_s2d_bf16: unsupported geometry B=%d H=%d W=%d (even H, W)", B, H, W);
    return CB2_ERR_UNSUPPORTED;
  }
  const long long total = static_cast<long long>(B) * (H / 2 + 3) * (W / 2 + 3);
  long long blocks = (total + 255) / 256;
  const long long cap = static_cast<long long>(sm_count()) * 16;
  if (blocks > cap) blocks = cap;
  u8_to_s2d_bf16_kernel<<<static_cast<int>(blocks), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      in, reinterpret_cast<__nv_bfloat16*>(out), B, H, W, mean3[0], mean3[1], mean3[2], inv_std3[0],
      inv_std3[1], inv_std3[2]);
  return check_cuda(cudaGetLastError(), "u8_to_s2d_bf16 launch");
}

// ---------------------------------------------------------------------------------------------
// Every s-th pixel of an NHWC tensor (the input of a strided 1x1 convolution) and its adjoint.
// 16-byte vectors; the adjoint writes the whole gradient in one pass (zeros where no pixel was read)
// instead of a memset followed by a strided copy.
// ---------------------------------------------------------------------------------------------
namespace cb2 {
__global__ void __launch_bounds__(256)
subsample_fwd_kernel(const uint4* __restrict__ x, uint4* __restrict__ y, int H, int W, int C8, int OH,
                     int OW, int s, size_t total) {
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>(i % C8);
    size_t p = i / C8;
    const int ow = static_cast<int>(p % OW);
    p /= OW;
    const int oh = static_cast<int>(p % OH);
    const size_t b = p / OH;
    y[i] = ld_stream_v4(x + ((b * H + static_cast<size_t>(oh) * s) * W + static_cast<size_t>(ow) * s) * C8 + c);
  }
}
__global__ void __launch_bounds__(256)
subsample_bwd_kernel(const uint4* __restrict__ g, uint4* __restrict__ dx, int H, int W, int C8, int OH,
                     int OW, int s, size_t total) {
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>(i % C8);
    size_t p = i / C8;
    const int w = static_cast<int>(p % W);
    p /= W;
    const int h = static_cast<int>(p % H);
    const size_t b = p / H;
    uint4 v = make_uint4(0, 0, 0, 0);
    if (h % s == 0 && w % s == 0)
      v = ld_stream_v4(g + ((b * OH + h / s) * OW + w / s) * C8 + c);
    dx[i] = v;
  }
}
}  // namespace cb2

extern "C" int cb2_subsample_nhwc(const void* x, void* y, int B, int H, int W, int C, int s, int backward,
                                  void* stream) {
  if (B <= 0 || H <= 0 || W <= 0 || C <= 0 || (C & 7) || s < 1 ||
      (reinterpret_cast<uintptr_t>(x) & 15) || (reinterpret_cast<uintptr_t>(y) & 15)) {
    cb2::set_last_error("subsample_nhwc: need C %% 8 == 0 and 16-byte aligned tensors (B=%d H=%d W=%d C=%d s=%d)",
                        B, H, W, C, s);
    return CB2_ERR_ARG;
  }
  const int OH = (H + s - 1) / s, OW = (W + s - 1) / s, C8 = C / 8;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (!backward) {
    const size_t total = static_cast<size_t>(B) * OH * OW * C8;
    const int grid = static_cast<int>(std::min<size_t>((total + 255) / 256, 148 * 16));
    cb2::subsample_fwd_kernel<<<grid, 256, 0, st>>>(reinterpret_cast<const uint4*>(x),
                                                    reinterpret_cast<uint4*>(y), H, W, C8, OH, OW, s, total);
  } else {
    // x = the gradient of the subsampled tensor [B, OH, OW, C], y = the full gradient [B, H, W, C]
    const size_t total = static_cast<size_t>(B) * H * W * C8;
    const int grid = static_cast<int>(std::min<size_t>((total + 255) / 256, 148 * 16));
    cb2::subsample_bwd_kernel<<<grid, 256, 0, st>>>(reinterpret_cast<const uint4*>(x),
                                                    reinterpret_cast<uint4*>(y), H, W, C8, OH, OW, s, total);
  }
  return cb2::check_cuda(cudaGetLastError(), "subsample_nhwc launch");
}

extern "C" int cb2_optim_tick(float* hyper, void* stream) {
  optim_tick_kernel<<<1, 32, 0, reinterpret_cast<cudaStream_t>(stream)>>>(hyper);
  return check_cuda(cudaGetLastError(), "optim_tick launch");
}

extern "C" int cb2_optim_flat(int opt, float* master, const void* grad, int grad_dtype, float* s1,
                              float* s2, void* param_out, int param_dtype, size_t n,
                              const float* hyper, int zero_grad, void* stream) {
  if (n == 0) return 0;
  if (!master && !param_out) {
    set_last_error("optim_flat: need master and/or param_out");
    return CB2_ERR_ARG;
  }
  if (!master && param_dtype != CB2_DTYPE_F32) {
    set_last_error("optim_flat: bf16 parameters need an fp32 master copy");
    return CB2_ERR_ARG;
  }
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  void* g = const_cast<void*>(grad);
  switch (opt) {
    case CB2_OPT_NONE: return launch_optim_flat<CB2_OPT_NONE>(master, g, grad_dtype, s1, s2, param_out, param_dtype, n, hyper, zero_grad, st);
    case CB2_OPT_ADAM_KERAS: return launch_optim_flat<CB2_OPT_ADAM_KERAS>(master, g, grad_dtype, s1, s2, param_out, param_dtype, n, hyper, zero_grad, st);
    case CB2_OPT_ADAM_TORCH: return launch_optim_flat<CB2_OPT_ADAM_TORCH>(master, g, grad_dtype, s1, s2, param_out, param_dtype, n, hyper, zero_grad, st);
    case CB2_OPT_ADAMW: return launch_optim_flat<CB2_OPT_ADAMW>(master, g, grad_dtype, s1, s2, param_out, param_dtype, n, hyper, zero_grad, st);
    case CB2_OPT_SGD: return launch_optim_flat<CB2_OPT_SGD>(master, g, grad_dtype, s1, s2, param_out, param_dtype, n, hyper, zero_grad, st);
    case CB2_OPT_RMSPROP: return launch_optim_flat<CB2_OPT_RMSPROP>(master, g, grad_dtype, s1, s2, param_out, param_dtype, n, hyper, zero_grad, st);
  }
  set_last_error("optim_flat: unknown optimizer %d", opt);
  return CB2_ERR_ARG;
}
