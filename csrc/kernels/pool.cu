// NHWC max pooling (forward with arg-max, backward by gather) for bf16 activations.
//
// torch's max_pool_backward_nhwc scatters with atomics and took 463 us for the ResNet stem
// (64 x 112 x 112 x 64 -> 56 x 56) against ~30 us of pure memory traffic; here the backward pass
// is a gather: every input pixel looks at the <= ceil(k/s)^2 windows that contain it and picks up
// dy where the saved arg-max (one byte per output element) points at it.  Threads own 8 channels
// (one 16-byte access), so warps read and write whole 128-byte lines.  Zero padding is implicit
// (Keras' ZeroPadding2D + 'valid' pooling, core/tests/examples/call_run_within_script_with_keras_fit.py
// via keras.applications.ResNet50: pool1_pad + pool1_pool).
#include "optim.cuh"

namespace cb2 {

struct PoolGeom {
  int N, H, W, C, OH, OW, k, s, pad;
};

// KS = k*16 + s for the compile-time specialisations (0: read k and s from the geometry).  All
// index arithmetic is 32-bit (the host rejects tensors with 2^31 or more 16-byte groups): 64-bit
// divisions made the first version of these kernels instruction-bound.
template <int KS>
__global__ void __launch_bounds__(256)
maxpool_fwd_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y,
                   uint8_t* __restrict__ idx, const PoolGeom g) {
  const int K = KS ? (KS >> 4) : g.k, S = KS ? (KS & 15) : g.s;
  const unsigned c8n = g.C >> 3;
  const unsigned total = static_cast<unsigned>(g.N) * g.OH * g.OW * c8n;
  for (unsigned t = blockIdx.x * blockDim.x + threadIdx.x; t < total; t += gridDim.x * blockDim.x) {
    const int c8 = static_cast<int>(t % c8n) * 8;
    unsigned r = t / c8n;
    const int ow = static_cast<int>(r % g.OW); r /= g.OW;
    const int oh = static_cast<int>(r % g.OH);
    const int n = static_cast<int>(r / g.OH);
    float best[8];
    int arg[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) { best[j] = -INFINITY; arg[j] = 0; }
#pragma unroll
    for (int kh = 0; kh < K; ++kh) {
      const int h = oh * S - g.pad + kh;
#pragma unroll
      for (int kw = 0; kw < K; ++kw) {
        const int w = ow * S - g.pad + kw;
        Vec8 v;
        if (h >= 0 && h < g.H && w >= 0 && w < g.W) {
          v = unpack8_bf16(*reinterpret_cast<const uint4*>(
              x + ((static_cast<size_t>(n) * g.H + h) * g.W + w) * g.C + c8));
        } else {
#pragma unroll
          for (int j = 0; j < 8; ++j) v.v[j] = 0.f;     // zero padding
        }
        const int code = kh * K + kw;
#pragma unroll
        for (int j = 0; j < 8; ++j)
          if (v.v[j] > best[j]) { best[j] = v.v[j]; arg[j] = code; }
      }
    }
    Vec8 o;
#pragma unroll
    for (int j = 0; j < 8; ++j) o.v[j] = best[j];
    const size_t oi = ((static_cast<size_t>(n) * g.OH + oh) * g.OW + ow) * g.C + c8;
    *reinterpret_cast<uint4*>(y + oi) = pack8_bf16(o);
    uint2 packed;
    packed.x = arg[0] | (arg[1] << 8) | (arg[2] << 16) | (arg[3] << 24);
    packed.y = arg[4] | (arg[5] << 8) | (arg[6] << 16) | (arg[7] << 24);
    *reinterpret_cast<uint2*>(idx + oi) = packed;
  }
}

template <int KS>
__global__ void __launch_bounds__(256)
maxpool_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const uint8_t* __restrict__ idx,
                   __nv_bfloat16* __restrict__ dx, const PoolGeom g) {
  const int K = KS ? (KS >> 4) : g.k, S = KS ? (KS & 15) : g.s;
  const unsigned c8n = g.C >> 3;
  const unsigned total = static_cast<unsigned>(g.N) * g.H * g.W * c8n;
  for (unsigned t = blockIdx.x * blockDim.x + threadIdx.x; t < total; t += gridDim.x * blockDim.x) {
    const int c8 = static_cast<int>(t % c8n) * 8;
    unsigned r = t / c8n;
    const int w = static_cast<int>(r % g.W); r /= g.W;
    const int h = static_cast<int>(r % g.H);
    const int n = static_cast<int>(r / g.H);
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = 0.f;
    // windows (oh, ow) with oh*s - pad <= h <= oh*s - pad + k - 1
    const int hp = h + g.pad, wp = w + g.pad;
    int oh0 = (hp - K + S) / S;          // ceil((hp - k + 1) / s)
    if (hp - K + 1 <= 0) oh0 = 0;
    int ow0 = (wp - K + S) / S;
    if (wp - K + 1 <= 0) ow0 = 0;
    const int oh1 = min(hp / S, g.OH - 1), ow1 = min(wp / S, g.OW - 1);
    for (int oh = oh0; oh <= oh1; ++oh) {
      const int kh = hp - oh * S;
      for (int ow = ow0; ow <= ow1; ++ow) {
        const int code = kh * K + (wp - ow * S);
        const size_t oi = ((static_cast<size_t>(n) * g.OH + oh) * g.OW + ow) * g.C + c8;
        const uint2 pk = *reinterpret_cast<const uint2*>(idx + oi);
        const Vec8 gy = unpack8_bf16(*reinterpret_cast<const uint4*>(dy + oi));
        // byte-wise compare of four arg-max codes at once: 0xff where the window's max is here
        const uint32_t code4 = static_cast<uint32_t>(code) * 0x01010101u;
        const uint32_t m0 = __vcmpeq4(pk.x, code4), m1 = __vcmpeq4(pk.y, code4);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const uint32_t mk = j < 4 ? m0 : m1;
          if (mk & (0x80u << ((j & 3) * 8))) acc[j] += gy.v[j];
        }
      }
    }
    Vec8 o;
#pragma unroll
    for (int j = 0; j < 8; ++j) o.v[j] = acc[j];
    st_stream_v4(dx + ((static_cast<size_t>(n) * g.H + h) * g.W + w) * g.C + c8, pack8_bf16(o));
  }
}

static int pool_grid(size_t total) {
  int sms = 148, dev = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  size_t want = (total + 255) / 256;
  const size_t cap = static_cast<size_t>(sms) * 16;
  if (want > cap) want = cap;
  if (want < 1) want = 1;
  return static_cast<int>(want);
}

static int pool_check(const PoolGeom& g) {
  if (static_cast<long long>(g.N) * g.H * g.W * (g.C >> 3) >= (1LL << 31)) {
    set_last_error("maxpool: tensor too large for 32-bit indexing");
    return CB2_ERR_UNSUPPORTED;
  }
  if (g.N <= 0 || g.H <= 0 || g.W <= 0 || g.C <= 0 || (g.C & 7) || g.k <= 0 || g.k > 15 ||
      g.s <= 0 || g.pad < 0 || g.pad >= g.k || g.OH != (g.H + 2 * g.pad - g.k) / g.s + 1 ||
      g.OW != (g.W + 2 * g.pad - g.k) / g.s + 1) {
    set_last_error("maxpool: unsupported geometry N=%d H=%d W=%d C=%d k=%d s=%d pad=%d OH=%d OW=%d",
                   g.N, g.H, g.W, g.C, g.k, g.s, g.pad, g.OH, g.OW);
    return CB2_ERR_UNSUPPORTED;
  }
  return 0;
}

}  // namespace cb2

using namespace cb2;

extern "C" int cb2_maxpool_nhwc_forward(const void* x, void* y, void* argmax, int N, int H, int W,
                                        int C, int OH, int OW, int k, int s, int pad, void* stream) {
  PoolGeom g{N, H, W, C, OH, OW, k, s, pad};
  int rc = pool_check(g);
  if (rc) return rc;
  const size_t total = static_cast<size_t>(N) * OH * OW * (C >> 3);
  auto kern = (k == 3 && s == 2) ? maxpool_fwd_kernel<3 * 16 + 2>
              : (k == 2 && s == 2) ? maxpool_fwd_kernel<2 * 16 + 2> : maxpool_fwd_kernel<0>;
  kern<<<pool_grid(total), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const __nv_bfloat16*>(x), reinterpret_cast<__nv_bfloat16*>(y),
      reinterpret_cast<uint8_t*>(argmax), g);
  return check_cuda(cudaGetLastError(), "maxpool_fwd launch");
}

extern "C" int cb2_maxpool_nhwc_backward(const void* dy, const void* argmax, void* dx, int N, int H,
                                         int W, int C, int OH, int OW, int k, int s, int pad,
                                         void* stream) {
  PoolGeom g{N, H, W, C, OH, OW, k, s, pad};
  int rc = pool_check(g);
  if (rc) return rc;
  const size_t total = static_cast<size_t>(N) * H * W * (C >> 3);
  auto kern = (k == 3 && s == 2) ? maxpool_bwd_kernel<3 * 16 + 2>
              : (k == 2 && s == 2) ? maxpool_bwd_kernel<2 * 16 + 2> : maxpool_bwd_kernel<0>;
  kern<<<pool_grid(total), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const __nv_bfloat16*>(dy), reinterpret_cast<const uint8_t*>(argmax),
      reinterpret_cast<__nv_bfloat16*>(dx), g);
  return check_cuda(cudaGetLastError(), "maxpool_bwd launch");
}
