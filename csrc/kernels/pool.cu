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
// divisions made the first version of these kernels instruction-bound.  The arithmetic is packed
// bf16x2 (max / compare-mask / masked add on two channels per instruction): max is exact in bf16,
// and a gradient element is the sum of at most ceil(k/s)^2 bf16 values, nearly always one.
__device__ __forceinline__ __nv_bfloat162 as_bf162(uint32_t w) {
  return *reinterpret_cast<__nv_bfloat162*>(&w);
}
__device__ __forceinline__ uint32_t as_u32(__nv_bfloat162 v) {
  return *reinterpret_cast<uint32_t*>(&v);
}

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
    uint32_t best[4], arg[4];          // two channels per word: bf16 values / 16-bit tap codes
#pragma unroll
    for (int j = 0; j < 4; ++j) { best[j] = 0xff80ff80u; arg[j] = 0; }     // -inf, -inf
    const __nv_bfloat16* xn = x + static_cast<size_t>(n) * g.H * g.W * g.C + c8;
#pragma unroll
    for (int kh = 0; kh < K; ++kh) {
      const int h = oh * S - g.pad + kh;
#pragma unroll
      for (int kw = 0; kw < K; ++kw) {
        const int w = ow * S - g.pad + kw;
        uint4 v = make_uint4(0, 0, 0, 0);             // zero padding
        if (h >= 0 && h < g.H && w >= 0 && w < g.W)
          v = *reinterpret_cast<const uint4*>(xn + (static_cast<size_t>(h) * g.W + w) * g.C);
        const uint32_t code2 = static_cast<uint32_t>(kh * K + kw) * 0x00010001u;
        const uint32_t vv[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const uint32_t m = __hgt2_mask(as_bf162(vv[j]), as_bf162(best[j]));   // strict: first max wins
          best[j] = (vv[j] & m) | (best[j] & ~m);
          arg[j] = (code2 & m) | (arg[j] & ~m);
        }
      }
    }
    const size_t oi = ((static_cast<size_t>(n) * g.OH + oh) * g.OW + ow) * g.C + c8;
    *reinterpret_cast<uint4*>(y + oi) = make_uint4(best[0], best[1], best[2], best[3]);
    uint2 packed;                                      // one byte per channel
    packed.x = __byte_perm(arg[0], arg[1], 0x6420);
    packed.y = __byte_perm(arg[2], arg[3], 0x6420);
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
    // specialised (k, s): at most four windows contain a pixel - packed bf16 accumulation;
    // generic: up to k*k contributions, accumulated in fp32
    __nv_bfloat162 acc[4];
    float accf[8];
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[j] = as_bf162(0u);
#pragma unroll
    for (int j = 0; j < 8; ++j) accf[j] = 0.f;
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
        const uint32_t code4 = static_cast<uint32_t>(kh * K + (wp - ow * S)) * 0x01010101u;
        const size_t oi = ((static_cast<size_t>(n) * g.OH + oh) * g.OW + ow) * g.C + c8;
        const uint2 pk = *reinterpret_cast<const uint2*>(idx + oi);
        const uint4 gy = *reinterpret_cast<const uint4*>(dy + oi);
        // byte-wise compare of four arg-max codes at once (0xff where this pixel was the max),
        // widened to 16-bit lane masks
        const uint32_t m0 = __vcmpeq4(pk.x, code4), m1 = __vcmpeq4(pk.y, code4);
        const uint32_t g0 = gy.x & __byte_perm(m0, 0, 0x1100), g1 = gy.y & __byte_perm(m0, 0, 0x3322);
        const uint32_t g2 = gy.z & __byte_perm(m1, 0, 0x1100), g3 = gy.w & __byte_perm(m1, 0, 0x3322);
        if (KS != 0) {
          acc[0] = __hadd2(acc[0], as_bf162(g0));
          acc[1] = __hadd2(acc[1], as_bf162(g1));
          acc[2] = __hadd2(acc[2], as_bf162(g2));
          acc[3] = __hadd2(acc[3], as_bf162(g3));
        } else {
          const uint32_t gg[4] = {g0, g1, g2, g3};
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            accf[2 * j] += __uint_as_float(gg[j] << 16);
            accf[2 * j + 1] += __uint_as_float(gg[j] & 0xffff0000u);
          }
        }
      }
    }
    uint4 o;
    if (KS != 0) {
      o = make_uint4(as_u32(acc[0]), as_u32(acc[1]), as_u32(acc[2]), as_u32(acc[3]));
    } else {
      o = make_uint4(pack_bf16(accf[0], accf[1]), pack_bf16(accf[2], accf[3]),
                     pack_bf16(accf[4], accf[5]), pack_bf16(accf[6], accf[7]));
    }
    st_stream_v4(dx + ((static_cast<size_t>(n) * g.H + h) * g.W + w) * g.C + c8, o);
  }
}

// Backward of the ResNet stem pool (k = 3, s = 2, pad = 1, even H and W): one thread per 2x2 block
// of input pixels and 8 channels.  The four pixels of a block lie in the same four windows
// (oh in {r, r+1}, ow in {s, s+1}), so their arg-max bytes and gradients are loaded once - the
// per-pixel gather above fetches 2.25 windows per pixel and was bound by that L1/L2 traffic
// (345 us at batch 256 for 0.56 GB of compulsory traffic).
__global__ void __launch_bounds__(256)
maxpool_bwd_3x3s2_kernel(const __nv_bfloat16* __restrict__ dy, const uint8_t* __restrict__ idx,
                         __nv_bfloat16* __restrict__ dx, const PoolGeom g) {
  const unsigned c8n = g.C >> 3;
  const int RH = g.H >> 1, RW = g.W >> 1;
  const unsigned total = static_cast<unsigned>(g.N) * RH * RW * c8n;
  for (unsigned t = blockIdx.x * blockDim.x + threadIdx.x; t < total; t += gridDim.x * blockDim.x) {
    const int c8 = static_cast<int>(t % c8n) * 8;
    unsigned q = t / c8n;
    const int s = static_cast<int>(q % RW); q /= RW;
    const int r = static_cast<int>(q % RH);
    const int n = static_cast<int>(q / RH);
    uint4 gy[2][2];
    uint2 pk[2][2];
#pragma unroll
    for (int a = 0; a < 2; ++a)
#pragma unroll
      for (int b = 0; b < 2; ++b) {
        gy[a][b] = make_uint4(0, 0, 0, 0);
        pk[a][b] = make_uint2(0xffffffffu, 0xffffffffu);          // matches no tap code
        if (r + a < g.OH && s + b < g.OW) {
          const size_t oi = ((static_cast<size_t>(n) * g.OH + r + a) * g.OW + s + b) * g.C + c8;
          pk[a][b] = *reinterpret_cast<const uint2*>(idx + oi);
          gy[a][b] = *reinterpret_cast<const uint4*>(dy + oi);
        }
      }
#pragma unroll
    for (int dh = 0; dh < 2; ++dh)
#pragma unroll
      for (int dw = 0; dw < 2; ++dw) {
        __nv_bfloat162 acc[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[j] = as_bf162(0u);
#pragma unroll
        for (int a = 0; a <= dh; ++a)
#pragma unroll
          for (int b = 0; b <= dw; ++b) {
            // tap of pixel (2r+dh, 2s+dw) inside window (r+a, s+b): kh = dh + 1 - 2a, kw = dw + 1 - 2b
            const uint32_t code4 = static_cast<uint32_t>((dh + 1 - 2 * a) * 3 + (dw + 1 - 2 * b)) * 0x01010101u;
            const uint32_t m0 = __vcmpeq4(pk[a][b].x, code4), m1 = __vcmpeq4(pk[a][b].y, code4);
            acc[0] = __hadd2(acc[0], as_bf162(gy[a][b].x & __byte_perm(m0, 0, 0x1100)));
            acc[1] = __hadd2(acc[1], as_bf162(gy[a][b].y & __byte_perm(m0, 0, 0x3322)));
            acc[2] = __hadd2(acc[2], as_bf162(gy[a][b].z & __byte_perm(m1, 0, 0x1100)));
            acc[3] = __hadd2(acc[3], as_bf162(gy[a][b].w & __byte_perm(m1, 0, 0x3322)));
          }
        st_stream_v4(dx + ((static_cast<size_t>(n) * g.H + 2 * r + dh) * g.W + 2 * s + dw) * g.C + c8,
                     make_uint4(as_u32(acc[0]), as_u32(acc[1]), as_u32(acc[2]), as_u32(acc[3])));
      }
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
  if (k == 3 && s == 2 && pad == 1 && !(H & 1) && !(W & 1)) {
    maxpool_bwd_3x3s2_kernel<<<pool_grid(total / 4), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
        reinterpret_cast<const __nv_bfloat16*>(dy), reinterpret_cast<const uint8_t*>(argmax),
        reinterpret_cast<__nv_bfloat16*>(dx), g);
    return check_cuda(cudaGetLastError(), "maxpool_bwd launch");
  }
  auto kern = (k == 3 && s == 2) ? maxpool_bwd_kernel<3 * 16 + 2>
              : (k == 2 && s == 2) ? maxpool_bwd_kernel<2 * 16 + 2> : maxpool_bwd_kernel<0>;
  kern<<<pool_grid(total), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const __nv_bfloat16*>(dy), reinterpret_cast<const uint8_t*>(argmax),
      reinterpret_cast<__nv_bfloat16*>(dx), g);
  return check_cuda(cudaGetLastError(), "maxpool_bwd launch");
}
