// 3x3 / stride-1 / pad-1 convolution (forward and input gradient) as an implicit GEMM on tcgen05.
//
//   y[b, h, w, co] = sum_{dy, dx, ci} x[b, h + dy - 1, w + dx - 1, ci] * k[co, dy, dx, ci]      (NHWC, bf16)
//
// No im2col and no per-tap reload of the activation.  One TMA load fetches, for a block of 64 input
// channels, a STRIP of the image: RT + 2 rows by W + 2 columns (the halo comes back as zeros from
// TMA's out-of-bounds fill), laid out pixel by pixel in 128-byte swizzled rows - exactly a K-major
// UMMA operand whose "rows" are the pixels of the zero-padded image in raster order.  In that
// order a tap (dy, dx) is a constant shift of dy (W + 2) + dx rows, so the nine taps are nine
// operand descriptors into the SAME strip: the strip is read from L2 once and multiplied nine times.
// The output tile is 256 padded-raster pixels (two 128-row MMA sub-tiles, each a whole number of
// image rows) by BN channels; the two "pixels" per row that fall into the padding are computed and
// thrown away (the TMA store of a tile is a box W + 2 wide over a tensor W wide: they are clipped).
//
// The two sub-tiles share every weight tile (each B stage feeds two MMAs), which halves the
// shared-memory-fill traffic per flop: with 128 rows per CTA the weights alone need 64 B/cycle/SM
// from L2, more than the chip delivers to 148 SMs at once.
//
//   warps: 0 strip producer, 1 MMA issuer + TMEM owner, 2 weight producer, 3 idle, 4-11 epilogue
//          (warp = sub-tile * 4 + TMEM lane quarter)
//   TMEM : 2 buffers x 2 sub-tiles x BN fp32 columns (BN = 128: all 512 columns)
//   input gradient: the same kernel on dy with the taps mirrored and the weight tile read
//          channel-transposed (MN-major B operand) - no transposed copy of the weights.
//   optional epilogue: per-channel sum / sum of squares of the stored values (the statistics of the
//          batch-norm that follows), only over real pixels.
//
// Workload: conv*_block*_2_conv of ResNet-50 (core/tests/examples/
// call_run_within_script_with_keras_fit.py:79-83); the reference leaves it to cuDNN inside TensorFlow.
#include <stdlib.h>

#include "gemm_host.cuh"

namespace cb2 {

constexpr int kC3Threads = 384;
constexpr int kC3MaxAStages = 4;
constexpr int kC3MaxBStages = 12;
constexpr int kC3SmemBudget = 232448 - 1024;  // dynamic shared memory, less the alignment slack
constexpr int kC3StatCopies = 8;

struct C3Args {
  int B, H, W;
  int Nc;                         // channels produced
  int tap_stride;                 // columns between two taps of the stored [Cout, 9 Cin] weight matrix
  int Wp, Rsub, RT;               // W + 2; image rows per sub-tile / per tile
  int tiles_per_image, m_tiles, n_tiles;
  int k_blocks;                   // reduction channels / 64
  int mirror;                     // input gradient: tap t uses weight tap 8 - t
  int base_off;                   // operand descriptors carry the swizzle phase of their first row
  int a_stage, a_stages;          // bytes of one strip stage (multiple of 1024); how many
  int b_stages;
  int out_bufs;                   // staging buffers per sub-tile (1 or 2)
  int full_units, total_units;    // work units: whole tiles first, then the last tiles as single sub-tiles
  int skip;                       // bring-up: bit 0 no strip loads, bit 1 no weight loads, bit 2 no stores
  float* colstats;                // nullable [kC3StatCopies][2][Nc]
};

struct alignas(64) C3Maps {
  CUtensorMap a, b, c;
};

template <int BN>
struct C3Cfg {
  static constexpr int kBStage = BN * 128;
  static constexpr int kTmemCols = 4 * BN;                 // 2 buffers x 2 sub-tiles
  static_assert(kTmemCols <= 512, "TMEM budget");
};

// K-major, 128B-swizzled operand whose first row sits at any 128-byte boundary of a swizzled region:
// the start address carries the row offset, and the base-offset field tells the MMA unit where in
// the 8-row swizzle period that row lies.
__device__ __forceinline__ uint64_t make_smem_desc_row(uint32_t saddr, int base_off) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((saddr >> 4) & 0x3FFFu);
  d |= static_cast<uint64_t>(1) << 16;                     // LBO (unused for K-major swizzled)
  d |= static_cast<uint64_t>(1024 >> 4) << 32;             // SBO: 8 rows
  d |= 1ull << 46;
  if (base_off) d |= static_cast<uint64_t>((saddr >> 7) & 7u) << 49;
  d |= 2ull << 61;                                         // SWIZZLE_128B
  return d;
}

// A work unit: a tile (two sub-tiles) or, at the tail of the schedule, one sub-tile of a tile - the
// last partial wave is cut in halves so it costs half a wave.  A sub-tile that starts past the last
// image row does no MMAs and stores nothing.
struct C3Unit {
  int b, h0, n0;
  int on;                         // bit s: sub-tile s has rows of the image
};
template <int BN>
__device__ __forceinline__ C3Unit c3_unit(const C3Args& p, int u) {
  int tile = u, half = -1;
  if (u >= p.full_units) {
    const int x = u - p.full_units;
    tile = p.full_units + (x >> 1);
    half = x & 1;
  } else {
    // Whole waves are rotated by one CTA per wave.  A tile whose second sub-tile lies past the image
    // costs half a tile, and such tiles come with a fixed period (28 x 28: every 4th); with 148 CTAs
    // - a multiple of 4 - striding through them, a quarter of the CTAs would get nothing but the
    // cheap tiles and the rest nothing but whole ones.
    const int G = gridDim.x, wave = u / G;
    if ((wave + 1) * G <= p.full_units) tile = wave * G + (u - wave * G + wave) % G;
  }
  const int mt = tile % p.m_tiles;
  C3Unit r;
  r.n0 = (tile / p.m_tiles) * BN;
  r.b = mt / p.tiles_per_image;
  r.h0 = (mt % p.tiles_per_image) * p.RT + (half == 1 ? p.Rsub : 0);
  r.on = (r.h0 < p.H ? 1 : 0) | ((half < 0 && r.h0 + p.Rsub < p.H) ? 2 : 0);
  return r;
}

template <int BN, bool B_MN>
__global__ void __launch_bounds__(kC3Threads, 1)
conv3x3_kernel(const __grid_constant__ C3Maps tm, const C3Args p) {
  using Cfg = C3Cfg<BN>;
  constexpr int kNC = BN / 64;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  const int kAS = p.a_stages, kBS = p.b_stages;
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem_a + kAS * p.a_stage;
  uint8_t* smem_out = smem_b + kBS * Cfg::kBStage;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_out + 2 * p.out_bufs * 16384);
  uint64_t* afull = bars;
  uint64_t* aempty = afull + kC3MaxAStages;
  uint64_t* bfull = aempty + kC3MaxAStages;
  uint64_t* bempty = bfull + kC3MaxBStages;
  uint64_t* tfull = bempty + kC3MaxBStages;     // [accumulator buffer][sub-tile]
  uint64_t* tempty = tfull + 4;                 // [accumulator buffer][sub-tile]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 4);

  // broadcast from lane 0: tells ptxas the value is warp-uniform, so what derives from it (stage
  // indices, descriptors) can live in uniform registers - where tcgen05 instructions want them
  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tm.a);
    tma_prefetch_desc(&tm.b);
    tma_prefetch_desc(&tm.c);
    // two MMA issuers (one per sub-tile) release every operand stage
    for (int s = 0; s < kAS; ++s) { mbar_init(&afull[s], 1); mbar_init(&aempty[s], 2); }
    for (int s = 0; s < kBS; ++s) { mbar_init(&bfull[s], 1); mbar_init(&bempty[s], 2); }
    for (int a = 0; a < 4; ++a) { mbar_init(&tfull[a], 1); mbar_init(&tempty[a], 4); }
    mbar_fence_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, Cfg::kTmemCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const int strip_bytes = (p.RT + 2) * p.Wp * 128;

  if (warp == 0) {
    // ===================== strips of the input =====================
    if (lane == 0) {
      int seq = 0;
      for (int u = blockIdx.x; u < p.total_units; u += gridDim.x) {
        const C3Unit un = c3_unit<BN>(p, u);
        const int b = un.b, h0 = un.h0;
        for (int cb = 0; cb < p.k_blocks; ++cb, ++seq) {
          const int st = seq % kAS;
          mbar_wait(&aempty[st], ((seq / kAS) & 1) ^ 1);
          if (p.skip & 1) { mbar_arrive(&afull[st]); continue; }
          mbar_expect_tx(&afull[st], strip_bytes);
          tma_load_4d(&tm.a, &afull[st], smem_a + st * p.a_stage, cb * 64, -1, h0 - 1, b);
        }
      }
    }
    __syncwarp();
  } else if (warp == 2) {
    // ===================== weight tiles: one per (channel block, tap) =====================
    if (lane == 0) {
      int seq = 0;
      for (int u = blockIdx.x; u < p.total_units; u += gridDim.x) {
        const int n0 = c3_unit<BN>(p, u).n0;
        for (int cb = 0; cb < p.k_blocks; ++cb)
          for (int tap = 0; tap < 9; ++tap, ++seq) {
            const int st = seq % kBS;
            const int wt = p.mirror ? 8 - tap : tap;
            mbar_wait(&bempty[st], ((seq / kBS) & 1) ^ 1);
            if (p.skip & 2) { mbar_arrive(&bfull[st]); continue; }
            mbar_expect_tx(&bfull[st], Cfg::kBStage);
            uint8_t* sb = smem_b + st * Cfg::kBStage;
            if (!B_MN) {
              // forward: rows = output channels, K = input channels of this tap
              tma_load_2d(&tm.b, &bfull[st], sb, wt * p.tap_stride + cb * 64, n0);
            } else {
              // input gradient: N = input channels (contiguous), K = output channels (rows)
#pragma unroll
              for (int c = 0; c < BN / 64; ++c)
                tma_load_2d(&tm.b, &bfull[st], sb + c * (64 * 128), wt * p.tap_stride + n0 + c * 64, cb * 64);
            }
          }
      }
    }
    __syncwarp();
  } else if (warp == 1 || warp == 3) {
    // ===================== MMA issuers: warp 1 sub-tile 0, warp 3 sub-tile 1 =====================
    // One thread cannot issue the short MMAs of this kernel (64 cycles each at BN = 128, 32 at
    // BN = 64) back to back - about ten dependent instructions sit between two of them - so each
    // sub-tile has its own issuer on its own scheduler; the tensor core interleaves the two streams.
    {
      const int s = warp >> 1;
      const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((B_MN ? 1u : 0u) << 16) |
                             (static_cast<uint32_t>(BN >> 3) << 17) | (static_cast<uint32_t>(128 >> 4) << 24);
      const uint32_t b_lbo = B_MN ? (64 * 128) >> 4 : 1, b_sbo = 1024 >> 4;
      const uint32_t b_kadv = B_MN ? (16 * 128) >> 4 : (16 * 2) >> 4;
      // descriptors: the high word is fixed, the low word is the operand's address in 16-byte units.
      // Every offset below is a handful of integer adds - the issue loop has to stay well under the
      // 8 x 64 cycles the tensor core needs for the eight MMAs of one tap.
      const uint64_t a_hi = make_smem_desc_row(0, 0) & 0xffffffff00000000ull;
      const uint64_t b_hi = make_smem_desc(0, b_lbo, b_sbo) & 0xffffffff00000000ull;
      const uint32_t lbo_lo = B_MN ? (b_lbo << 16) : (1u << 16);
      const uint32_t a_base = smem_u32(smem_a) >> 4, b_base = smem_u32(smem_b) >> 4;
      const uint32_t a_step = static_cast<uint32_t>(p.a_stage) >> 4, b_step = Cfg::kBStage >> 4;
      const uint32_t wp8 = static_cast<uint32_t>(p.Wp) * 8;              // one image row of the strip
      const uint32_t sub8 = static_cast<uint32_t>(p.Rsub) * wp8;         // second sub-tile
      int ast = 0, bst = 0, t = 0;
      uint32_t aph = 0, bph = 0;
      for (int u = blockIdx.x; u < p.total_units; u += gridDim.x) {
        if (!((c3_unit<BN>(p, u).on >> s) & 1)) {
          // nothing to compute: hand the operand stages back as they arrive
          for (int cb = 0; cb < p.k_blocks; ++cb) {
            mbar_wait_warp(&afull[ast], aph);
            for (int tap = 0; tap < 9; ++tap) {
              mbar_wait_warp(&bfull[bst], bph);
              if (lane == 0) mbar_arrive(&bempty[bst]);
              if (++bst == kBS) { bst = 0; bph ^= 1; }
            }
            if (lane == 0) mbar_arrive(&aempty[ast]);
            if (++ast == kAS) { ast = 0; aph ^= 1; }
          }
          continue;
        }
        const int acc = t & 1;
        mbar_wait_warp(&tempty[acc * 2 + s], ((t >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t d = tmem_base + acc * (2 * BN) + s * BN;
        for (int cb = 0; cb < p.k_blocks; ++cb) {
          mbar_wait_warp(&afull[ast], aph);
          tc_fence_after();
          const uint32_t strip = a_base + ast * a_step;
#pragma unroll
          for (int tap = 0; tap < 9; ++tap) {
            mbar_wait_warp(&bfull[bst], bph);
            tc_fence_after();
            const uint32_t b_lo = ((b_base + bst * b_step) & 0x3FFFu) | lbo_lo;
            const uint32_t a_lo = (strip + s * sub8 + (tap / 3) * wp8 + (tap % 3) * 8) | (1u << 16);
#pragma unroll
            for (int k = 0; k < 4; ++k)
              umma_bf16_warp(d, a_hi | (a_lo + k * 2), b_hi | (b_lo + k * b_kadv), idesc, (cb | tap | k) != 0);
            umma_commit_warp(&bempty[bst]);
            if (++bst == kBS) { bst = 0; bph ^= 1; }
          }
          umma_commit_warp(&aempty[ast]);
          if (++ast == kAS) { ast = 0; aph ^= 1; }
        }
        umma_commit_warp(&tfull[acc * 2 + s]);
        ++t;
      }
    }
    __syncwarp();
  } else if (warp >= 4) {
    // ===================== epilogue =====================
    const int q = warp & 3;
    const int s = (warp - 4) >> 2;
    const uint32_t sw = lane & 7;
    const int r = q * 32 + lane;                        // padded-raster pixel inside the sub-tile
    float cs[kNC][4];
#pragma unroll
    for (int j = 0; j < kNC; ++j) cs[j][0] = cs[j][1] = cs[j][2] = cs[j][3] = 0.f;
    int cs_n0 = -1;
    int buf = 0, t = 0;
    const int bar_id = 1 + s;
    for (int u = blockIdx.x; u < p.total_units; u += gridDim.x) {
      const C3Unit un = c3_unit<BN>(p, u);
      if (!((un.on >> s) & 1)) continue;
      const int acc = t & 1;
      const int n0 = un.n0, b = un.b, h0 = un.h0 + s * p.Rsub;
      // which of this warp's 32 rows are pixels of the image (not padding columns, not past the last row)
      const int hh = r / p.Wp, ww = r - hh * p.Wp;
      const uint32_t okmask = __ballot_sync(0xffffffffu, hh < p.Rsub && ww < p.W && h0 + hh < p.H);
      if (p.colstats != nullptr && cs_n0 != n0) {
        if (cs_n0 >= 0) {
          float* dst = p.colstats + static_cast<size_t>((blockIdx.x * 8 + warp) % kC3StatCopies) * 2 * p.Nc;
#pragma unroll
          for (int j = 0; j < kNC; ++j) {
            const int n = cs_n0 + j * 64 + 2 * lane;
            atomicAdd(dst + n, cs[j][0]);
            atomicAdd(dst + n + 1, cs[j][1]);
            atomicAdd(dst + p.Nc + n, cs[j][2]);
            atomicAdd(dst + p.Nc + n + 1, cs[j][3]);
            cs[j][0] = cs[j][1] = cs[j][2] = cs[j][3] = 0.f;
          }
        }
        cs_n0 = n0;
      }
      mbar_wait(&tfull[acc * 2 + s], (t >> 1) & 1);
      tc_fence_after();
      if (p.skip & 8) {
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&tempty[acc * 2 + s]);
        ++t;
        continue;
      }
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * (2 * BN) + s * BN;
#pragma unroll
      for (int j = 0; j < kNC; ++j) {
        uint32_t v[2][32];
        tmem_ld_32x32(taddr + j * 64, v[0]);
        tmem_ld_32x32(taddr + j * 64 + 32, v[1]);
        const uint32_t stage = smem_u32(smem_out) + (s * p.out_bufs + buf) * 16384;
        const uint32_t rowp = stage + r * 128;
        if (q == 0 && lane == 0) {                              // the store that last read this buffer is done
          if (p.out_bufs == 2) tma_store_wait_read<1>();
          else tma_store_wait_read<0>();
        }
        asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");
        tmem_ld_wait_for(v[0]);
        reg_fence(v[1]);
#pragma unroll
        for (int h = 0; h < 2; ++h)
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            const uint32_t* w = &v[h][g * 8];
            st_shared_v4(rowp + (((h * 4 + g) ^ sw) << 4), pack_bf16(__uint_as_float(w[0]), __uint_as_float(w[1])),
                         pack_bf16(__uint_as_float(w[2]), __uint_as_float(w[3])),
                         pack_bf16(__uint_as_float(w[4]), __uint_as_float(w[5])),
                         pack_bf16(__uint_as_float(w[6]), __uint_as_float(w[7])));
          }
        fence_proxy_async_smem();
        asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");
        if (q == 0 && lane == 0 && !(p.skip & 4)) {
          // box {64 channels, W + 2, Rsub rows}: the two padding columns and rows past H are clipped
          tma_store_4d(&tm.c, stage, n0 + j * 64, 0, h0, b);
          tma_store_commit();
        }
        if (p.colstats != nullptr) {
          // column sums of the staged bf16 values over the real pixels of this warp's 32 rows:
          // lane l owns channels 2l, 2l+1 of the chunk and walks down the rows
          const uint32_t colp = stage + q * 32 * 128 + ((lane & 3) << 2);
          const uint32_t jj = lane >> 2;
#pragma unroll
          for (int r0 = 0; r0 < 32; r0 += 16) {
            uint32_t wv[16];
#pragma unroll
            for (int rr = 0; rr < 16; ++rr)
              asm volatile("ld.shared.b32 %0, [%1];"
                           : "=r"(wv[rr])
                           : "r"(colp + (r0 + rr) * 128 + ((jj ^ ((r0 + rr) & 7)) << 4)));
#pragma unroll
            for (int rr = 0; rr < 16; ++rr) {
              const bool ok = (okmask >> (r0 + rr)) & 1u;
              const float a = ok ? __uint_as_float(wv[rr] << 16) : 0.f;
              const float bb = ok ? __uint_as_float(wv[rr] & 0xffff0000u) : 0.f;
              cs[j][0] += a;
              cs[j][1] += bb;
              cs[j][2] = fmaf(a, a, cs[j][2]);
              cs[j][3] = fmaf(bb, bb, cs[j][3]);
            }
          }
        }
        buf = (buf + 1) & (p.out_bufs - 1);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty[acc * 2 + s]);
      ++t;
    }
    if (p.colstats != nullptr && cs_n0 >= 0) {
      float* dst = p.colstats + static_cast<size_t>((blockIdx.x * 8 + warp) % kC3StatCopies) * 2 * p.Nc;
#pragma unroll
      for (int j = 0; j < kNC; ++j) {
        const int n = cs_n0 + j * 64 + 2 * lane;
        atomicAdd(dst + n, cs[j][0]);
        atomicAdd(dst + n + 1, cs[j][1]);
        atomicAdd(dst + p.Nc + n, cs[j][2]);
        atomicAdd(dst + p.Nc + n + 1, cs[j][3]);
      }
    }
    if (q == 0 && lane == 0) tma_store_wait<0>();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, Cfg::kTmemCols);
}

template <int BN, bool B_MN>
static int launch_c3(const C3Maps& maps, const C3Args& args, cudaStream_t stream) {
  auto kern = conv3x3_kernel<BN, B_MN>;
  static bool configured = false;
  if (!configured) {
    CB2_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kC3SmemBudget + 1024));
    configured = true;
  }
  const int grid = args.total_units < num_sms() ? args.total_units : num_sms();
  kern<<<grid, kC3Threads, kC3SmemBudget + 1024, stream>>>(maps, args);
  return check_cuda(cudaGetLastError(), "conv3x3 launch");
}

static int c3_tmap_4d(CUtensorMap* tm, const void* base, int C, int W, int H, int B, int box_c, int box_w,
                      int box_h) {
  const cuuint64_t dims[4] = {static_cast<cuuint64_t>(C), static_cast<cuuint64_t>(W),
                              static_cast<cuuint64_t>(H), static_cast<cuuint64_t>(B)};
  const cuuint64_t strides[3] = {static_cast<cuuint64_t>(C) * 2, static_cast<cuuint64_t>(W) * C * 2,
                                 static_cast<cuuint64_t>(H) * W * C * 2};
  const cuuint32_t box[4] = {static_cast<cuuint32_t>(box_c), static_cast<cuuint32_t>(box_w),
                             static_cast<cuuint32_t>(box_h), 1};
  return make_tmap_nd(tm, base, 4, dims, strides, box);
}

}  // namespace cb2

using namespace cb2;

static int g_c3_base_off = 0, g_c3_a_stages = 0, g_c3_b_stages = 0, g_c3_out_bufs = 0, g_c3_skip = 0;
// bring-up / tuning switches (0 = the default): strip descriptors with a base offset (wrong on sm_100a -
// the swizzle phase comes from the address itself), pipeline depths, staging buffers, skipped traffic
extern "C" int cb2_conv3x3_debug(int base_off, int a_stages, int b_stages, int out_bufs, int skip) {
  g_c3_base_off = base_off;
  g_c3_a_stages = a_stages;
  g_c3_b_stages = b_stages;
  g_c3_out_bufs = out_bufs;
  g_c3_skip = skip;
  return 0;
}

extern "C" int cb2_conv3x3_supported(int H, int W, int Cin, int Cout) {
  return H >= 1 && W >= 6 && W <= 62 && Cin % 64 == 0 && Cout % 64 == 0 && Cin >= 64 && Cout >= 64;
}

extern "C" int cb2_conv3x3(const void* x, const void* w, void* y, int B, int H, int W, int Cin, int Cout,
                           int input_gradient, float* colstats, void* stream) {
  // forward: x [B,H,W,Cin], w [Cout,3,3,Cin], y [B,H,W,Cout].
  // input gradient: x = dy [B,H,W,Cout], w the same tensor, y = dx [B,H,W,Cin].
  const int K = input_gradient ? Cout : Cin;      // reduction channels
  const int N = input_gradient ? Cin : Cout;      // produced channels
  if (!cb2_conv3x3_supported(H, W, Cin, Cout) || B <= 0 || (reinterpret_cast<uintptr_t>(x) & 15) ||
      (reinterpret_cast<uintptr_t>(w) & 15) || (reinterpret_cast<uintptr_t>(y) & 15)) {
    set_last_error("conv3x3: unsupported geometry B=%d H=%d W=%d Cin=%d Cout=%d (6 <= W <= 62, channels "
                   "multiples of 64, 16-byte aligned dense NHWC tensors)", B, H, W, Cin, Cout);
    return CB2_ERR_UNSUPPORTED;
  }
  C3Args a;
  a.B = B; a.H = H; a.W = W;
  a.Nc = N;
  a.tap_stride = Cin;
  a.Wp = W + 2;
  a.Rsub = 128 / a.Wp;
  if (a.Rsub > H) a.Rsub = H;                      // small images: one sub-tile holds them
  a.RT = 2 * a.Rsub;
  a.tiles_per_image = (H + a.RT - 1) / a.RT;
  a.m_tiles = B * a.tiles_per_image;
  a.k_blocks = K / 64;
  a.mirror = input_gradient ? 1 : 0;
  a.colstats = colstats;
  a.base_off = g_c3_base_off;
  const int bn = N % 128 == 0 ? 128 : 64;
  if (N % bn) {
    set_last_error("conv3x3: produced channels (%d) must be a multiple of the tile width %d", N, bn);
    return CB2_ERR_UNSUPPORTED;
  }
  a.n_tiles = N / bn;
  // shared memory: strip stages (the last tap of the second sub-tile reads 128 rows from row
  // (Rsub + 2) Wp + 2: the stage covers that), weight stages with what is left
  const int last_row = (a.Rsub + 2) * a.Wp + 2 + 128;
  a.a_stage = (last_row * 128 + 1023) & ~1023;
  // one staging buffer per sub-tile: the epilogue has slack (the kernel is MMA-bound), the weight
  // pipeline does not - the 32 KiB buy two more stages
  a.out_bufs = g_c3_out_bufs ? g_c3_out_bufs : 1;
  a.a_stages = g_c3_a_stages ? g_c3_a_stages : 2;
  if (a.a_stages > kC3MaxAStages) a.a_stages = kC3MaxAStages;
  const int fixed = 2 * a.out_bufs * 16384 + 512;
  int left = kC3SmemBudget - fixed - a.a_stages * a.a_stage;
  a.b_stages = left / (bn * 128);
  if (a.b_stages > kC3MaxBStages) a.b_stages = kC3MaxBStages;
  if (g_c3_b_stages && g_c3_b_stages < a.b_stages) a.b_stages = g_c3_b_stages;
  if (a.b_stages < 2) {
    set_last_error("conv3x3: no room for the weight pipeline (strip stage %d bytes x %d)", a.a_stage, a.a_stages);
    return CB2_ERR_UNSUPPORTED;
  }
  a.skip = g_c3_skip;
  {
    // the last partial wave, when it fills less than half of the SMs, runs as single sub-tiles
    const int tiles = a.m_tiles * a.n_tiles, sms = num_sms();
    const int rem = tiles > sms ? tiles % sms : 0;
    a.full_units = (rem > 0 && 2 * rem <= sms && !(g_c3_skip & 16)) ? tiles - rem : tiles;
    a.total_units = a.full_units + 2 * (tiles - a.full_units);
  }
  C3Maps maps;
  int rc;
  if ((rc = c3_tmap_4d(&maps.a, x, K, W, H, B, 64, a.Wp, a.RT + 2))) return rc;
  if ((rc = c3_tmap_4d(&maps.c, y, N, W, H, B, 64, a.Wp, a.Rsub))) return rc;
  // the weights as a matrix [Cout rows, 9 Cin columns]; forward reads [bn rows, 64 columns] tiles
  // (K-major B), the input gradient [64 rows, 64 columns] tiles (MN-major B)
  const uint64_t cols = static_cast<uint64_t>(9) * Cin;
  if ((rc = make_tmap(&maps.b, w, cols, Cout, cols, 64, input_gradient ? 64 : bn))) return rc;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (!input_gradient) return bn == 128 ? launch_c3<128, false>(maps, a, st) : launch_c3<64, false>(maps, a, st);
  return bn == 128 ? launch_c3<128, true>(maps, a, st) : launch_c3<64, true>(maps, a, st);
}
