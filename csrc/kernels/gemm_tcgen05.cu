// Dense-layer GEMM for B200 (sm_100a): C[M,N] = op(A)[M,K] * op(B)[N,K]^T with a fused epilogue.
//
//   * operands: bf16, staged global->shared by TMA (cp.async.bulk.tensor, 128B swizzle),
//   * math: tcgen05.mma (kind::f16, cta_group::1, 128 x BN x 16) issued by one thread,
//   * accumulators: fp32 in TMEM, double buffered (2 x BN columns) so the epilogue of tile i
//     overlaps the main loop of tile i+1,
//   * persistent: one CTA per SM, static round-robin tile schedule,
//   * warp roles: warp0 = TMA producer, warp1 = TMEM owner + MMA issuer, warps2-5 = epilogue,
//   * epilogue: tcgen05.ld (next chunk prefetched) -> bias / ReLU -> 128B-swizzled smem tile ->
//     TMA store, or TMA reduce-add for ACCUM / split-K; optional per-column sum / sum of squares
//     (batch-norm statistics); a per-row path with direct stores remains for the row-softmax and
//     ReLU-mask epilogues,
//   * split-K: work item = (m tile, n tile, K slice); slices meet in an fp32 C through reduce-adds,
//     optionally rounded into a bf16 matrix by the last-arriving warp (cb2_gemm_splitk_bf16),
//   * convolutions as implicit GEMMs: the producer can address NHWC activations in place through
//     3-D / 4-D tensor maps (cb2_conv_stem_forward / _wgrad, cb2_conv_wgrad at the end of the file).
//
// Both operands may be K-major (reduction dim contiguous) or MN-major, which covers the three
// GEMMs of a Dense layer without any transpose kernel:
//   forward  Y  = X  W^T        A=X  [M,K]  K-major,  B=W  [N,K] K-major
//   dgrad    dX = dY W          A=dY [M,N'] K-major,  B=W  stored [N',K'] -> MN-major
//   wgrad    dW = dY^T X        A=dY stored [Mb,N] -> MN-major, B=X stored [Mb,K] -> MN-major
//
// Reference parity: this is op K1/K2/K3 of SURVEY.md section 2.5 (the Dense layers the reference's
// workloads run through TensorFlow, e.g. README.md:71-75); the reference itself has no kernel.
#include <cuda.h>
#include <stdlib.h>

#include <mutex>

#include "common.cuh"
#include "gemm_host.cuh"
#include "../../include/cloud_b200_kernels.h"

namespace cb2 {

constexpr int BM = 128;
constexpr int BK = 64;          // 64 bf16 = one 128-byte swizzle span
constexpr int UMMA_K = 16;
constexpr int kGemmThreads = 192;
constexpr int kAStageBytes = BM * BK * 2;

struct GemmArgs {
  int M, N, K;
  int ldc, ldmask;
  void* C;
  const float* bias;
  const __nv_bfloat16* mask;
  uint32_t flags;
  int m_tiles, n_tiles, k_blocks;
  uint32_t a_lbo, a_sbo, b_lbo, b_sbo;  // descriptor fields, 16-byte units
  uint32_t a_kadv, b_kadv;              // start-address advance per UMMA_K step, 16-byte units
  int tma_store;                        // epilogue through smem staging + TMA store (coalesced)
  int debug;                            // bench/gemm_epilogue_probe.py: skip parts of the epilogue
  int splits, kb_per_split;             // split-K: work item = (m tile, n tile, k slice)
  int m_fast;                           // work order: m tile index fastest (else n fastest)
  // ResNet stem as an implicit GEMM over the space-to-depth input (see cb2_conv_stem_*):
  // 1 = forward (one m tile per output row, 4 k blocks = the 4 kernel rows, A through a 4-D map,
  // C through a 3-D map that clips the 16 unused rows of the 128-row tile); 2 = weight gradient
  // (k block = 64 output pixels of one output row, A/B through 4-D / 3-D maps)
  int conv, conv_oh, conv_a_bytes;
  // 3 = weight gradient of a k x k / stride-1 convolution: GEMM rows = (tap, input channel), GEMM
  // K = output pixels in boxes of 64 (w_box x h_box); the A chunk of a tap is the same pixel box
  // shifted by the tap offset, out-of-image pixels zero-filled by TMA
  int conv_cin, conv_taps, conv_kw, conv_pad, conv_hblocks, conv_hbox, conv_stride;
  float* colstats;                      // [kColStatCopies][2][N] += column sums / sums of squares
  __nv_bfloat16* fin_out;               // split-K finalize: fin_out += bf16(C) by the last slice
  int fin_ld;                           //   row pitch of fin_out (elements)
  unsigned* fin_counters;               //   [m_tiles * n_tiles * 4], zero between calls
};

template <int BN>
struct GemmCfg {
  static constexpr int kStages = BN == 64 ? 8 : (BN == 128 ? 6 : 4);
  static constexpr int kBStageBytes = BN * BK * 2;
  static constexpr int kStageBytes = kAStageBytes + kBStageBytes;
  static constexpr int kTmemCols = 2 * BN;
  // epilogue staging: 4 warps x 2 buffers x (32 rows x 128 B), 128B-swizzled, feeds TMA stores
  static constexpr int kStagingBytes = 4 * 2 * 4096 + BN * 4;   // + bias copy
  static_assert(kStages * kStageBytes + kStagingBytes + 1024 + 256 <= 232448, "smem budget");
  static constexpr int kSmemBytes =
      kStages * kStageBytes + kStagingBytes + 1024 /*align slack*/ + 256 /*barriers*/;
};

// Work item -> (m tile, n tile, k-block range).  Consecutive items differ in n (default: CTAs that
// run together share the A tile) or in m (m_fast: they share the B tile, and one CTA keeps the same
// n for long runs - what the column-statistics epilogue wants).
struct WorkItem { int m0, n0, kb0, kb1; };
template <int BN>
__device__ __forceinline__ WorkItem decode_work(const GemmArgs& p, int w) {
  int mi, ni, si;
  if (p.m_fast) {
    mi = w % p.m_tiles; w /= p.m_tiles;
    ni = w % p.n_tiles; si = w / p.n_tiles;
  } else {
    ni = w % p.n_tiles; w /= p.n_tiles;
    mi = w % p.m_tiles; si = w / p.m_tiles;
  }
  WorkItem it;
  it.m0 = mi * BM;
  it.n0 = ni * BN;
  it.kb0 = si * p.kb_per_split;
  it.kb1 = min(p.k_blocks, it.kb0 + p.kb_per_split);
  return it;
}

constexpr int kColStatCopies = 8;   // same replica count as the batch-norm scratch (bn.cu)

// Coalesced epilogue of one 32-row slice of an accumulator tile: TMEM -> registers -> 128B-swizzled
// smem tile (32 rows x 128 B, double buffered per warp) -> TMA store / reduce-add.  Rows and
// columns outside the matrix are clipped by the tensor map.  Each scheduler runs a single
// epilogue warp, so every latency in here is exposed: the TMEM load of chunk i+1 is issued before
// chunk i is converted, the element loop is straight-line code (bias from a smem copy, ReLU as a
// max with 0 or -inf, or neither when `plain`), and all addresses are explicit shared addresses.
template <int BN, bool F32, bool STATS>
__device__ __forceinline__ void epilogue_tma(const CUtensorMap* tmC, uint32_t taddr, uint32_t stg,
                                             uint32_t sbias, int& stage_buf, int n0, int row0, int N,
                                             float lo, bool plain, bool accum, int lane,
                                             float (&cs)[F32 ? 1 : BN / 64][4], int conv_row = -1) {
  // unit = 32 accumulator columns = one tcgen05.ld; a staged chunk (128-byte rows) is one unit of
  // fp32 or two units of bf16
  constexpr int NU = BN / 32;
  constexpr int UPC = F32 ? 1 : 2;
  uint32_t v[2][32];
  const uint32_t sw = lane & 7;
  const int rem = N - n0;
  const int nu = rem >= BN ? NU : (((rem + 32 * UPC - 1) / (32 * UPC)) * UPC);
  tmem_ld_32x32(taddr, v[0]);
#pragma unroll
  for (int u = 0; u < NU; ++u) {
    if (u < nu) {
      uint32_t(&cur)[32] = v[u & 1];
      tmem_ld_wait_for(cur);
      if (u + 1 < NU && u + 1 < nu) tmem_ld_32x32(taddr + (u + 1) * 32, v[(u + 1) & 1]);
      const int ci = u / UPC, h = u % UPC;
      const uint32_t buf = stg + stage_buf * 4096;
      const uint32_t rowp = buf + lane * 128;
      if (h == 0) {
        if (lane == 0) tma_store_wait_read<1>();   // the store that last read this buffer is done
        __syncwarp();
      }
      if (F32) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          if (plain) {
            st_shared_v4(rowp + ((j ^ sw) << 4), cur[j * 4], cur[j * 4 + 1], cur[j * 4 + 2], cur[j * 4 + 3]);
          } else {
            const float4 bb = ld_shared_f4(sbias + (u * 32 + j * 4) * 4);
            const float f0 = fmaxf(__uint_as_float(cur[j * 4 + 0]) + bb.x, lo);
            const float f1 = fmaxf(__uint_as_float(cur[j * 4 + 1]) + bb.y, lo);
            const float f2 = fmaxf(__uint_as_float(cur[j * 4 + 2]) + bb.z, lo);
            const float f3 = fmaxf(__uint_as_float(cur[j * 4 + 3]) + bb.w, lo);
            st_shared_v4(rowp + ((j ^ sw) << 4), __float_as_uint(f0), __float_as_uint(f1),
                         __float_as_uint(f2), __float_as_uint(f3));
          }
        }
      } else {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const uint32_t* w = &cur[j * 8];
          uint32_t o0, o1, o2, o3;
          if (plain) {
            o0 = pack_bf16(__uint_as_float(w[0]), __uint_as_float(w[1]));
            o1 = pack_bf16(__uint_as_float(w[2]), __uint_as_float(w[3]));
            o2 = pack_bf16(__uint_as_float(w[4]), __uint_as_float(w[5]));
            o3 = pack_bf16(__uint_as_float(w[6]), __uint_as_float(w[7]));
          } else {
            const float4 b0 = ld_shared_f4(sbias + (u * 32 + j * 8) * 4);
            const float4 b1 = ld_shared_f4(sbias + (u * 32 + j * 8 + 4) * 4);
            o0 = pack_bf16(fmaxf(__uint_as_float(w[0]) + b0.x, lo), fmaxf(__uint_as_float(w[1]) + b0.y, lo));
            o1 = pack_bf16(fmaxf(__uint_as_float(w[2]) + b0.z, lo), fmaxf(__uint_as_float(w[3]) + b0.w, lo));
            o2 = pack_bf16(fmaxf(__uint_as_float(w[4]) + b1.x, lo), fmaxf(__uint_as_float(w[5]) + b1.y, lo));
            o3 = pack_bf16(fmaxf(__uint_as_float(w[6]) + b1.z, lo), fmaxf(__uint_as_float(w[7]) + b1.w, lo));
          }
          st_shared_v4(rowp + (((h * 4 + j) ^ sw) << 4), o0, o1, o2, o3);
        }
      }
      if (h == UPC - 1) {
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) {
          if (conv_row >= 0) tma_store_3d(tmC, buf, n0 + ci * (32 * UPC), row0, conv_row);
          else if (accum) tma_reduce_add_2d(tmC, buf, n0 + ci * (32 * UPC), row0);
          else tma_store_2d(tmC, buf, n0 + ci * (32 * UPC), row0);
          tma_store_commit();
        }
        if (STATS && !F32) {
          // Column sums of what was just staged (the bf16-rounded values the next kernel will
          // read): lane l owns columns 2l, 2l+1 of the chunk and walks down the 32 rows; with
          // the 128B swizzle the 32 lanes of one row read hit 32 different banks.
          const uint32_t colp = buf + ((lane & 3) << 2);
          const uint32_t j = lane >> 2;
          // loads in batches of 16 ahead of the arithmetic: with one warp per scheduler nothing
          // else hides the shared-memory latency
#pragma unroll
          for (int r0 = 0; r0 < 32; r0 += 16) {
            uint32_t wv[16];
#pragma unroll
            for (int r = 0; r < 16; ++r)
              asm volatile("ld.shared.b32 %0, [%1];"
                           : "=r"(wv[r])
                           : "r"(colp + (r0 + r) * 128 + ((j ^ ((r0 + r) & 7)) << 4)));
            float s0 = 0.f, s1 = 0.f, q0 = 0.f, q1 = 0.f;      // second set of chains: more ILP
#pragma unroll
            for (int r = 0; r < 16; r += 2) {
              const float a = __uint_as_float(wv[r] << 16), b = __uint_as_float(wv[r] & 0xffff0000u);
              const float c = __uint_as_float(wv[r + 1] << 16), d = __uint_as_float(wv[r + 1] & 0xffff0000u);
              cs[ci][0] += a;
              cs[ci][1] += b;
              cs[ci][2] = fmaf(a, a, cs[ci][2]);
              cs[ci][3] = fmaf(b, b, cs[ci][3]);
              s0 += c;
              s1 += d;
              q0 = fmaf(c, c, q0);
              q1 = fmaf(d, d, q1);
            }
            cs[ci][0] += s0;
            cs[ci][1] += s1;
            cs[ci][2] += q0;
            cs[ci][3] += q1;
          }
        }
        stage_buf ^= 1;
      }
    }
  }
}

// Add a warp's column statistics of tile column block n0 to the global accumulator and clear them.
template <int BN>
__device__ __forceinline__ void flush_colstats(float (&cs)[BN / 64][4], float* __restrict__ out,
                                               int n0, int N, int replica, int lane) {
  float* dst = out + static_cast<size_t>(replica) * 2 * N;
#pragma unroll
  for (int ci = 0; ci < BN / 64; ++ci) {
    const int n = n0 + ci * 64 + 2 * lane;
    if (n < N) {
      atomicAdd(dst + n, cs[ci][0]);
      atomicAdd(dst + N + n, cs[ci][2]);
    }
    if (n + 1 < N) {
      atomicAdd(dst + n + 1, cs[ci][1]);
      atomicAdd(dst + N + n + 1, cs[ci][3]);
    }
    cs[ci][0] = cs[ci][1] = cs[ci][2] = cs[ci][3] = 0.f;
  }
}

// Split-K finalize, done by the kernel itself.  Every epilogue warp that has added its K slice of
// a 32-row output slice into the fp32 workspace C bumps that slice's counter; the warp that
// arrives last (counter == splits - 1) owns the complete sum: it adds it, rounded to bf16, into
// fin_out (the gradient arena), puts the workspace slice back to zero and resets the counter, so
// neither a reduction kernel nor a per-step memset of the workspace is needed.
template <int BN>
__device__ __forceinline__ void splitk_finalize(const GemmArgs& p, int row0, int n0, int slot,
                                                int lane) {
  unsigned old = 0;
  if (lane == 0) {
    tma_store_wait<0>();          // this warp's reduce-adds have been performed at L2
    __threadfence();
    old = atomicAdd(p.fin_counters + slot, 1u);
  }
  old = __shfl_sync(0xffffffffu, old, 0);
  if (old != static_cast<unsigned>(p.splits - 1)) return;
  __threadfence();
  constexpr int CPL = BN / 32;                 // columns per lane: 2, 4 or 8
  constexpr int RB = 8;                        // rows per batch: all their loads are issued first
  const int c = n0 + lane * CPL;
  float* ws = reinterpret_cast<float*>(p.C);
  if (c < p.N) {
#pragma unroll 1
    for (int r0 = 0; r0 < 32; r0 += RB) {
      float v[RB][CPL];
      uint32_t o[RB][CPL / 2];
#pragma unroll
      for (int r = 0; r < RB; ++r) {
        const int row = min(row0 + r0 + r, p.M - 1);          // clamped: loads stay in bounds
        const float* wp = ws + static_cast<size_t>(row) * p.ldc + c;
        const __nv_bfloat16* op = p.fin_out + static_cast<size_t>(row) * p.fin_ld + c;
        if (CPL == 2) {
          const float2 t = __ldcg(reinterpret_cast<const float2*>(wp));
          v[r][0] = t.x; v[r][1] = t.y;
          o[r][0] = *reinterpret_cast<const uint32_t*>(op);
        } else if (CPL == 4) {
          const float4 t = __ldcg(reinterpret_cast<const float4*>(wp));
          v[r][0] = t.x; v[r][1] = t.y; v[r][2] = t.z; v[r][3] = t.w;
          const uint2 u = *reinterpret_cast<const uint2*>(op);
          o[r][0] = u.x; o[r][CPL / 2 - 1] = u.y;
        } else {
          const float4 t0 = __ldcg(reinterpret_cast<const float4*>(wp));
          const float4 t1 = __ldcg(reinterpret_cast<const float4*>(wp + 4));
          v[r][0] = t0.x; v[r][1] = t0.y; v[r][2] = t0.z; v[r][3] = t0.w;
          v[r][CPL - 4] = t1.x; v[r][CPL - 3] = t1.y; v[r][CPL - 2] = t1.z; v[r][CPL - 1] = t1.w;
          const uint4 u = *reinterpret_cast<const uint4*>(op);
          o[r][0] = u.x; o[r][1 % (CPL / 2)] = u.y; o[r][CPL / 2 - 2 < 0 ? 0 : CPL / 2 - 2] = u.z;
          o[r][CPL / 2 - 1] = u.w;
        }
      }
#pragma unroll
      for (int r = 0; r < RB; ++r) {
        const int row = row0 + r0 + r;
        if (row < p.M) {
          float* wp = ws + static_cast<size_t>(row) * p.ldc + c;
          __nv_bfloat16* op = p.fin_out + static_cast<size_t>(row) * p.fin_ld + c;
#pragma unroll
          for (int k = 0; k < CPL / 2; ++k)
            o[r][k] = pack_bf16(__uint_as_float(o[r][k] << 16) + v[r][2 * k],
                                __uint_as_float(o[r][k] & 0xffff0000u) + v[r][2 * k + 1]);
          if (CPL == 2) {
            *reinterpret_cast<float2*>(wp) = make_float2(0.f, 0.f);
            *reinterpret_cast<uint32_t*>(op) = o[r][0];
          } else if (CPL == 4) {
            *reinterpret_cast<float4*>(wp) = make_float4(0.f, 0.f, 0.f, 0.f);
            *reinterpret_cast<uint2*>(op) = make_uint2(o[r][0], o[r][CPL / 2 - 1]);
          } else {
            *reinterpret_cast<float4*>(wp) = make_float4(0.f, 0.f, 0.f, 0.f);
            *reinterpret_cast<float4*>(wp + 4) = make_float4(0.f, 0.f, 0.f, 0.f);
            *reinterpret_cast<uint4*>(op) = make_uint4(o[r][0], o[r][1 % (CPL / 2)],
                                                       o[r][CPL / 2 - 2 < 0 ? 0 : CPL / 2 - 2],
                                                       o[r][CPL / 2 - 1]);
          }
        }
      }
    }
  }
  __syncwarp();
  if (lane == 0) p.fin_counters[slot] = 0;
}

template <int BN, bool A_MN, bool B_MN>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_tcgen05_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                    const __grid_constant__ CUtensorMap tmC, const GemmArgs p) {
  using Cfg = GemmCfg<BN>;
  constexpr int kStages = Cfg::kStages;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + kStages * kAStageBytes;
  uint8_t* smem_stage = smem + kStages * Cfg::kStageBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_stage + Cfg::kStagingBytes);
  uint64_t* full_bar = bars;                    // [kStages] TMA -> MMA
  uint64_t* empty_bar = bars + kStages;         // [kStages] MMA -> TMA
  uint64_t* tfull_bar = bars + 2 * kStages;     // [2] MMA -> epilogue
  uint64_t* tempty_bar = bars + 2 * kStages + 2;  // [2] epilogue -> MMA
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * kStages + 4);
  uint64_t* wbar = bars + 2 * kStages + 5;      // conv == 4: resident weights loaded

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);    // warp-uniform for ptxas
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    if (p.tma_store) tma_prefetch_desc(&tmC);
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tfull_bar[a], 1);
      mbar_init(&tempty_bar[a], 4);
    }
    mbar_init(wbar, 1);
    mbar_fence_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, Cfg::kTmemCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const int num_tiles = p.m_tiles * p.n_tiles * p.splits;
  // conv == 4 (stem forward with an input-row ring): every CTA owns a contiguous range of output
  // rows, so that consecutive tiles share three of their four input rows; otherwise round-robin
  // (the convolution modes exist only in the instantiations their host entry points launch)
  constexpr bool kStemFwd = BN == 64 && !A_MN && !B_MN;
  constexpr bool kConvMN = A_MN && B_MN;
  int tile_begin = blockIdx.x, tile_end = num_tiles, tile_step = gridDim.x;
  if (kStemFwd && p.conv == 4) {
    const int per = (num_tiles + gridDim.x - 1) / gridDim.x;
    tile_begin = blockIdx.x * per;
    tile_end = min(num_tiles, tile_begin + per);
    tile_step = 1;
  }

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      if (kStemFwd && p.conv == 4) {
        // resident weights: the four k blocks go into B stages 0..3 once
        mbar_expect_tx(wbar, 4 * Cfg::kBStageBytes);
        for (int kb = 0; kb < 4; ++kb)
          tma_load_2d(&tmB, wbar, smem_b + kb * Cfg::kBStageBytes, kb * BK, 0);
        // input rows stream through the A stages as a ring: tile (b, oh) needs rows oh..oh+3 and
        // only row oh+3 is new unless the tile starts an image or this CTA's range
        int seq = 0;
        for (int R = tile_begin; R < tile_end; ++R) {
          const int b = R / p.conv_oh, oh = R - b * p.conv_oh;
          const bool fresh = R == tile_begin || oh == 0;
          for (int h = fresh ? oh : oh + 3; h <= oh + 3; ++h, ++seq) {
            const int slot = seq % kStages;
            mbar_wait(&empty_bar[slot], (((seq / kStages) & 1) ^ 1));
            mbar_expect_tx(&full_bar[slot], p.conv_a_bytes);
            tma_load_4d(&tmA, &full_bar[slot], smem_a + slot * kAStageBytes, 0, 0, h, b);
          }
        }
        tile_begin = tile_end;       // nothing left for the generic loop below
      }
      for (int tile = tile_begin; tile < tile_end; tile += tile_step) {
        const WorkItem it = decode_work<BN>(p, tile);
        const int m0 = it.m0, n0 = it.n0;
        for (int kb = it.kb0; kb < it.kb1; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem_a + stage * kAStageBytes;
          uint8_t* sb = smem_b + stage * Cfg::kBStageBytes;
          const int k0 = kb * BK;
          if (kStemFwd && p.conv == 1) {
            // stem forward: output row R = (image, oh); k block kb = kernel row: the 112 A rows
            // are the 128-byte windows of input row oh + kb, one per output pixel (32 B apart)
            const int R = m0 / BM;
            mbar_expect_tx(&full_bar[stage], p.conv_a_bytes + Cfg::kBStageBytes);
            tma_load_4d(&tmA, &full_bar[stage], sa, 0, 0, R % p.conv_oh + kb, R / p.conv_oh);
            tma_load_2d(&tmB, &full_bar[stage], sb, k0, n0);
            if (++stage == kStages) { stage = 0; phase ^= 1; }
            continue;
          }
          mbar_expect_tx(&full_bar[stage], Cfg::kStageBytes);
          if (kConvMN && BN == 64 && p.conv == 2) {
            // stem weight gradient: k block = 64 output pixels (half) of output row R; the two
            // 64-wide chunks of the A tile are kernel rows 2*mt and 2*mt + 1
            const int R = kb >> 1, half = kb & 1;
#pragma unroll
            for (int c = 0; c < BM / 64; ++c)
              tma_load_4d(&tmA, &full_bar[stage], sa + c * (BK * 128), 0, half * 64,
                          R % p.conv_oh + (m0 / 64) + c, R / p.conv_oh);
            tma_load_3d(&tmB, &full_bar[stage], sb, 0, half * 64, R);
            if (++stage == kStages) { stage = 0; phase ^= 1; }
            continue;
          }
          if (kConvMN && p.conv == 3) {
            const int b = kb / p.conv_hblocks;
            const int h0 = (kb - b * p.conv_hblocks) * p.conv_hbox;
#pragma unroll
            for (int c = 0; c < BM / 64; ++c) {
              const int g = m0 + c * 64;
              const int tap = g / p.conv_cin;
              const int ci0 = g - tap * p.conv_cin;
              const int i = tap / p.conv_kw, j = tap - i * p.conv_kw;
              // rows past the last tap: a box that lies outside the image in h -> all zeros
              const int hh = tap < p.conv_taps ? h0 * p.conv_stride + i - p.conv_pad : -(1 << 20);
              tma_load_4d(&tmA, &full_bar[stage], sa + c * (BK * 128), ci0, j - p.conv_pad, hh, b);
            }
#pragma unroll
            for (int c = 0; c < BN / 64; ++c)
              tma_load_4d(&tmB, &full_bar[stage], sb + c * (BK * 128), n0 + c * 64, 0, h0, b);
            if (++stage == kStages) { stage = 0; phase ^= 1; }
            continue;
          }
          if (!A_MN) {
            tma_load_2d(&tmA, &full_bar[stage], sa, k0, m0);  // box {64 k, 128 rows}
          } else {
#pragma unroll
            for (int c = 0; c < BM / 64; ++c)  // box {64 m, 64 k}
              tma_load_2d(&tmA, &full_bar[stage], sa + c * (BK * 128), m0 + c * 64, k0);
          }
          if (!B_MN) {
            tma_load_2d(&tmB, &full_bar[stage], sb, k0, n0);  // box {64 k, BN rows}
          } else {
#pragma unroll
            for (int c = 0; c < BN / 64; ++c)  // box {64 n, 64 k}
              tma_load_2d(&tmB, &full_bar[stage], sb + c * (BK * 128), n0 + c * 64, k0);
          }
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ===================== MMA issuer (whole warp, one elected lane issues: common.cuh) ==========
    {
      const uint32_t idesc = (1u << 4) /*D=f32*/ | (1u << 7) /*A=bf16*/ | (1u << 10) /*B=bf16*/ |
                             ((A_MN ? 1u : 0u) << 15) | ((B_MN ? 1u : 0u) << 16) |
                             (static_cast<uint32_t>(BN >> 3) << 17) |
                             (static_cast<uint32_t>(BM >> 4) << 24);
      int stage = 0;
      uint32_t phase = 0;
      int t = 0;
      if (kStemFwd && p.conv == 4) {
        mbar_wait_warp(wbar, 0);
        int base = 0, loaded = 0;    // load-sequence number of row oh of the tile / rows requested so far
        for (int R = tile_begin; R < tile_end; ++R, ++t) {
          const int oh = R % p.conv_oh;
          const bool fresh = R == tile_begin || oh == 0;
          base = fresh ? loaded : base + 1;
          loaded += fresh ? 4 : 1;
          const int acc = t & 1;
          mbar_wait_warp(&tempty_bar[acc], ((t >> 1) & 1) ^ 1);
          tc_fence_after();
          const uint32_t tmem_d = tmem_base + acc * BN;
          for (int i = 0; i < 4; ++i) {
            const int sq = base + i, slot = sq % kStages;
            mbar_wait_warp(&full_bar[slot], (sq / kStages) & 1);     // rows seen by earlier tiles: already complete
            tc_fence_after();
            const uint64_t adesc =
                make_smem_desc(smem_u32(smem_a + slot * kAStageBytes), p.a_lbo, p.a_sbo);
            const uint64_t bdesc =
                make_smem_desc(smem_u32(smem_b + i * Cfg::kBStageBytes), p.b_lbo, p.b_sbo);
#pragma unroll
            for (int k = 0; k < BK / UMMA_K; ++k)
              umma_bf16_warp(tmem_d, adesc + static_cast<uint64_t>(k * p.a_kadv),
                        bdesc + static_cast<uint64_t>(k * p.b_kadv), idesc, (i | k) != 0);
          }
          // this tile is the last user of row oh; the last tile of an image / of the range also
          // releases the three rows nobody else will read
          umma_commit_warp(&empty_bar[base % kStages]);
          if (R + 1 == tile_end || oh == p.conv_oh - 1)
            for (int i = 1; i < 4; ++i) umma_commit_warp(&empty_bar[(base + i) % kStages]);
          umma_commit_warp(&tfull_bar[acc]);
        }
        tile_begin = tile_end;
      }
      for (int tile = tile_begin; tile < tile_end; tile += tile_step, ++t) {
        const int acc = t & 1;
        const uint32_t acc_phase = (t >> 1) & 1;
        mbar_wait_warp(&tempty_bar[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + acc * BN;
        const WorkItem it = decode_work<BN>(p, tile);
        for (int kb = it.kb0; kb < it.kb1; ++kb) {
          mbar_wait_warp(&full_bar[stage], phase);
          tc_fence_after();
          const uint64_t adesc =
              make_smem_desc(smem_u32(smem_a + stage * kAStageBytes), p.a_lbo, p.a_sbo);
          const uint64_t bdesc =
              make_smem_desc(smem_u32(smem_b + stage * Cfg::kBStageBytes), p.b_lbo, p.b_sbo);
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k) {
            umma_bf16_warp(tmem_d, adesc + static_cast<uint64_t>(k * p.a_kadv),
                      bdesc + static_cast<uint64_t>(k * p.b_kadv), idesc, ((kb - it.kb0) | k) != 0);
          }
          umma_commit_warp(&empty_bar[stage]);  // smem slot reusable once these MMAs retire
          if (kb == it.kb1 - 1) umma_commit_warp(&tfull_bar[acc]);
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
    __syncwarp();
  } else {
    // ===================== epilogue (4 warps, TMEM lane quarter = warp % 4) =====================
    const int q = warp & 3;
    const bool out_f32 = (p.flags & CB2_GEMM_OUT_F32) != 0;
    const bool accum = (p.flags & CB2_GEMM_ACCUM) != 0;
    const bool relu = (p.flags & CB2_GEMM_RELU) != 0;
    const bool softmax = (p.flags & CB2_GEMM_SOFTMAX) != 0;
    const bool has_bias = p.bias != nullptr;
    const bool has_mask = p.mask != nullptr;
    const bool vec_ok =
        out_f32 ? ((p.ldc & 3) == 0 && (reinterpret_cast<uintptr_t>(p.C) & 15) == 0)
                : ((p.ldc & 7) == 0 && (reinterpret_cast<uintptr_t>(p.C) & 15) == 0);
    int t = 0;
    int stage_buf = 0;
    float cs[BN / 64][4];            // column statistics carried across tiles with the same n0
    int cs_n0 = -1;
#pragma unroll
    for (int ci = 0; ci < BN / 64; ++ci) cs[ci][0] = cs[ci][1] = cs[ci][2] = cs[ci][3] = 0.f;
    for (int tile = tile_begin; tile < tile_end; tile += tile_step, ++t) {
      const int acc = t & 1;
      const uint32_t acc_phase = (t >> 1) & 1;
      const WorkItem it = decode_work<BN>(p, tile);
      const int m0 = it.m0, n0 = it.n0;
      const int row = m0 + q * 32 + lane;
      const bool row_ok = row < p.M;
      mbar_wait(&tfull_bar[acc], acc_phase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * BN;

      if (p.tma_store) {
        // one bias copy for the four epilogue warps (they are at most one tile apart, hence the
        // named barrier before it is overwritten and after it is filled)
        const uint32_t sbias = smem_u32(smem_stage) + 4 * 8192;
        if (has_bias || t == 0) {
          asm volatile("bar.sync 1, 128;" ::: "memory");
          for (int i = threadIdx.x - 64; i < BN; i += 128) {
            const int n = n0 + i;
            const float bv = (has_bias && n < p.N) ? __ldg(p.bias + n) : 0.f;
            asm volatile("st.shared.f32 [%0], %1;" ::"r"(sbias + i * 4), "f"(bv) : "memory");
          }
          asm volatile("bar.sync 1, 128;" ::: "memory");
        }
        const float lo = relu ? 0.f : -INFINITY;
        const bool plain = !has_bias && !relu;
        const uint32_t stg = smem_u32(smem_stage) + (warp - 2) * 8192;
        if (!(p.debug & 16)) {
          if (out_f32) {
            float none[1][4];
            epilogue_tma<BN, true, false>(&tmC, taddr, stg, sbias, stage_buf, n0, m0 + q * 32, p.N, lo,
                                          plain, accum, lane, none);
          } else if (p.colstats != nullptr) {
            if (n0 != cs_n0) {
              if (cs_n0 >= 0)
                flush_colstats<BN>(cs, p.colstats, cs_n0, p.N, (blockIdx.x * 4 + q) % kColStatCopies, lane);
              cs_n0 = n0;
            }
            epilogue_tma<BN, false, true>(&tmC, taddr, stg, sbias, stage_buf, n0, m0 + q * 32, p.N, lo,
                                          plain, accum, lane, cs);
          } else {
            if (kStemFwd && (p.conv == 1 || p.conv == 4))
              epilogue_tma<BN, false, false>(&tmC, taddr, stg, sbias, stage_buf, n0, q * 32, p.N, lo,
                                             plain, accum, lane, cs, m0 / BM);
            else
              epilogue_tma<BN, false, false>(&tmC, taddr, stg, sbias, stage_buf, n0, m0 + q * 32, p.N, lo,
                                             plain, accum, lane, cs);
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&tempty_bar[acc]);
        if (p.fin_out != nullptr)
          splitk_finalize<BN>(p, m0 + q * 32, n0, ((m0 / BM) * p.n_tiles + n0 / BN) * 4 + q, lane);
        continue;
      }

      float row_max = -INFINITY, row_sum = 0.f;
      if (softmax) {
        // N <= BN (checked on the host): the whole row lives in this thread's TMEM lane.
        for (int c0 = 0; c0 < BN && n0 + c0 < p.N; c0 += 32) {
          uint32_t v[32];
          tmem_ld_32x32(taddr + c0, v);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            const int n = n0 + c0 + i;
            if (n < p.N) {
              float f = __uint_as_float(v[i]) + (has_bias ? __ldg(p.bias + n) : 0.f);
              row_max = fmaxf(row_max, f);
            }
          }
        }
        for (int c0 = 0; c0 < BN && n0 + c0 < p.N; c0 += 32) {
          uint32_t v[32];
          tmem_ld_32x32(taddr + c0, v);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            const int n = n0 + c0 + i;
            if (n < p.N) {
              float f = __uint_as_float(v[i]) + (has_bias ? __ldg(p.bias + n) : 0.f);
              row_sum += exp2f((f - row_max) * 1.4426950408889634f);
            }
          }
        }
        row_sum = 1.f / row_sum;
      }

      for (int c0 = 0; c0 < BN && n0 + c0 < p.N; c0 += 32) {
        uint32_t v[32];
        tmem_ld_32x32(taddr + c0, v);
        tmem_ld_wait();
        float f[32];
        const int nb = n0 + c0;
        const bool full = nb + 32 <= p.N;
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          f[i] = __uint_as_float(v[i]);
          if (has_bias && (full || nb + i < p.N)) f[i] += __ldg(p.bias + nb + i);
          if (relu) f[i] = fmaxf(f[i], 0.f);
          if (softmax) f[i] = exp2f((f[i] - row_max) * 1.4426950408889634f) * row_sum;
        }
        if (row_ok) {
        if (has_mask) {
          const __nv_bfloat16* mrow = p.mask + static_cast<size_t>(row) * p.ldmask + nb;
#pragma unroll
          for (int i = 0; i < 32; ++i)
            if ((full || nb + i < p.N) && !(__bfloat162float(mrow[i]) > 0.f)) f[i] = 0.f;
        }
        if (out_f32) {
          float* crow = reinterpret_cast<float*>(p.C) + static_cast<size_t>(row) * p.ldc + nb;
          if (full && vec_ok) {
#pragma unroll
            for (int i = 0; i < 32; i += 4) {
              float4 o = make_float4(f[i], f[i + 1], f[i + 2], f[i + 3]);
              if (accum) {
                const float4 old = *reinterpret_cast<const float4*>(crow + i);
                o.x += old.x; o.y += old.y; o.z += old.z; o.w += old.w;
              }
              *reinterpret_cast<float4*>(crow + i) = o;
            }
          } else {
            for (int i = 0; i < 32; ++i)
              if (nb + i < p.N) crow[i] = accum ? crow[i] + f[i] : f[i];
          }
        } else {
          __nv_bfloat16* crow =
              reinterpret_cast<__nv_bfloat16*>(p.C) + static_cast<size_t>(row) * p.ldc + nb;
          if (full && vec_ok) {
#pragma unroll
            for (int i = 0; i < 32; i += 8) {
              uint4 o;
              o.x = pack_bf16(f[i], f[i + 1]);
              o.y = pack_bf16(f[i + 2], f[i + 3]);
              o.z = pack_bf16(f[i + 4], f[i + 5]);
              o.w = pack_bf16(f[i + 6], f[i + 7]);
              *reinterpret_cast<uint4*>(crow + i) = o;
            }
          } else {
            for (int i = 0; i < 32; ++i)
              if (nb + i < p.N) crow[i] = __float2bfloat16(f[i]);
          }
        }
        }  // row_ok
        __syncwarp();  // reconverge before the next warp-collective tcgen05.ld
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty_bar[acc]);
    }
    if (p.colstats != nullptr && cs_n0 >= 0)
      flush_colstats<BN>(cs, p.colstats, cs_n0, p.N, (blockIdx.x * 4 + q) % kColStatCopies, lane);
    if (p.tma_store && lane == 0) tma_store_wait<0>();   // staging smem must outlive its stores
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, Cfg::kTmemCols);
}

// ---------------------------------------------------------------------------------------------
// Host side
// ---------------------------------------------------------------------------------------------
struct DescOverride {
  bool on = false;
  uint32_t a_lbo, a_sbo, b_lbo, b_sbo, a_kadv, b_kadv;
};
static DescOverride g_desc_override;

template <int BN, bool A_MN, bool B_MN>
static int launch_gemm(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmC,
                       const GemmArgs& args, cudaStream_t stream) {
  using Cfg = GemmCfg<BN>;
  auto kern = gemm_tcgen05_kernel<BN, A_MN, B_MN>;
  static bool configured = false;
  if (!configured) {
    CB2_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     Cfg::kSmemBytes));
    configured = true;
  }
  const int tiles = args.m_tiles * args.n_tiles * args.splits;
  const int grid = tiles < num_sms() ? tiles : num_sms();
  kern<<<grid, kGemmThreads, Cfg::kSmemBytes, stream>>>(tmA, tmB, tmC, args);
  return check_cuda(cudaGetLastError(), "gemm_tcgen05 launch");
}

}  // namespace cb2

using namespace cb2;

static int g_epilogue_debug = 0;
extern "C" int cb2_gemm_debug_epilogue(int bits) {
  g_epilogue_debug = bits;
  return 0;
}

extern "C" int cb2_gemm_debug_descriptors(int enable, unsigned a_lbo, unsigned a_sbo, unsigned b_lbo,
                                          unsigned b_sbo, unsigned a_kadv, unsigned b_kadv) {
  g_desc_override.on = enable != 0;
  g_desc_override.a_lbo = a_lbo;
  g_desc_override.a_sbo = a_sbo;
  g_desc_override.b_lbo = b_lbo;
  g_desc_override.b_sbo = b_sbo;
  g_desc_override.a_kadv = a_kadv;
  g_desc_override.b_kadv = b_kadv;
  return 0;
}

extern "C" int cb2_gemm_bf16_ex(const void* A, const void* B, void* C, const float* bias,
                                const void* relu_mask, int M, int N, int K, int lda, int ldb,
                                int ldc, int ldmask, int a_mn, int b_mn, unsigned flags, int block_n,
                                int split_k, float* colstats, void* stream);

namespace {
struct FinalizeArgs {
  void* out = nullptr;
  int ld = 0;
  unsigned* counters = nullptr;
  int n_counters = 0;
};
thread_local FinalizeArgs g_finalize;
}  // namespace

extern "C" int cb2_gemm_splitk_bf16(const void* A, const void* B, float* workspace, void* out_bf16,
                                    unsigned* counters, int n_counters, int M, int N, int K, int lda,
                                    int ldb, int ld_ws, int ld_out, int a_mn, int b_mn, int block_n,
                                    int split_k, void* stream) {
  if ((reinterpret_cast<uintptr_t>(out_bf16) & 15) || (ld_out & 7) || (ld_ws & 3) || counters == nullptr) {
    set_last_error("gemm_splitk: out must be 16-byte aligned with a row pitch that is a multiple of 8");
    return CB2_ERR_ARG;
  }
  g_finalize.out = out_bf16;
  g_finalize.ld = ld_out;
  g_finalize.counters = counters;
  g_finalize.n_counters = n_counters;
  const int rc = cb2_gemm_bf16_ex(A, B, workspace, nullptr, nullptr, M, N, K, lda, ldb, ld_ws, 0, a_mn,
                                  b_mn, CB2_GEMM_OUT_F32 | CB2_GEMM_ACCUM, block_n, split_k, nullptr,
                                  stream);
  g_finalize = FinalizeArgs();
  return rc;
}

extern "C" int cb2_gemm_bf16(const void* A, const void* B, void* C, const float* bias,
                             const void* relu_mask, int M, int N, int K, int lda, int ldb, int ldc,
                             int ldmask, int a_mn, int b_mn, unsigned flags, int block_n,
                             void* stream) {
  return cb2_gemm_bf16_ex(A, B, C, bias, relu_mask, M, N, K, lda, ldb, ldc, ldmask, a_mn, b_mn, flags,
                          block_n, 1, nullptr, stream);
}

extern "C" int cb2_gemm_colstat_copies(void) { return kColStatCopies; }

extern "C" int cb2_gemm_bf16_ex(const void* A, const void* B, void* C, const float* bias,
                                const void* relu_mask, int M, int N, int K, int lda, int ldb,
                                int ldc, int ldmask, int a_mn, int b_mn, unsigned flags, int block_n,
                                int split_k, float* colstats, void* stream) {
  if (M <= 0 || N <= 0 || K <= 0) {
    set_last_error("gemm: empty problem M=%d N=%d K=%d", M, N, K);
    return CB2_ERR_ARG;
  }
  if ((reinterpret_cast<uintptr_t>(A) & 15) || (reinterpret_cast<uintptr_t>(B) & 15) || (lda & 7) ||
      (ldb & 7)) {
    set_last_error("gemm: TMA needs 16-byte aligned operand bases and row pitches (lda=%d ldb=%d)",
                   lda, ldb);
    return CB2_ERR_ARG;
  }
  if ((flags & CB2_GEMM_ACCUM) && !(flags & CB2_GEMM_OUT_F32)) {
    set_last_error("gemm: ACCUM requires OUT_F32");
    return CB2_ERR_ARG;
  }
  const bool softmax = (flags & CB2_GEMM_SOFTMAX) != 0;
  int bn = block_n;
  const bool want_split = split_k != 1 && (flags & CB2_GEMM_OUT_F32) && (flags & CB2_GEMM_ACCUM);
  if (bn == 0 && want_split) {
    // split-K fills the machine whatever the tile count is: take the widest tile (best MMA shape,
    // least operand traffic); measured on the ResNet-50 wgrad shapes (bench/wgrad_probe.py)
    bn = N >= 256 ? 256 : (N >= 128 ? 128 : 64);
  }
  if (bn == 0) {
    if (N <= 64) bn = 64;
    else if (N <= 128) bn = 128;
    else {
      const int mt = (M + BM - 1) / BM;
      const int sms = num_sms();
      if (mt * ((N + 255) / 256) >= sms) bn = 256;
      else if (mt * ((N + 127) / 128) >= sms / 2) bn = 128;
      else bn = 64;
      if (softmax) bn = 256;
    }
  }
  if (bn != 64 && bn != 128 && bn != 256) {
    set_last_error("gemm: block_n must be 64, 128 or 256 (got %d)", bn);
    return CB2_ERR_ARG;
  }
  if (softmax && N > bn) {
    set_last_error("gemm: softmax epilogue needs N <= block_n (N=%d)", N);
    return CB2_ERR_ARG;
  }

  CUtensorMap tmA, tmB;
  int rc;
  if (!a_mn) rc = make_tmap(&tmA, A, K, M, lda, BK, BM);
  else       rc = make_tmap(&tmA, A, M, K, lda, 64, BK);
  if (rc) return rc;
  if (!b_mn) rc = make_tmap(&tmB, B, K, N, ldb, BK, bn);
  else       rc = make_tmap(&tmB, B, N, K, ldb, 64, BK);
  if (rc) return rc;

  // Epilogue through TMA stores whenever the output is TMA-addressable and no per-element side
  // input (ReLU mask) or row-wide operation (softmax) is requested.
  const bool out_f32 = (flags & CB2_GEMM_OUT_F32) != 0;
  const int esize = out_f32 ? 4 : 2;
  static const bool tma_store_disabled = getenv("TFC_GEMM_NO_TMA_STORE") != nullptr;
  const bool tma_store = !tma_store_disabled && !softmax && relu_mask == nullptr &&
                         (reinterpret_cast<uintptr_t>(C) & 15) == 0 &&
                         ((static_cast<long long>(ldc) * esize) & 15) == 0;
  CUtensorMap tmC = tmA;
  if (tma_store) {
    rc = make_tmap(&tmC, C, N, M, ldc, out_f32 ? 32 : 64, 32, out_f32);
    if (rc) return rc;
  }

  GemmArgs args;
  args.tma_store = tma_store ? 1 : 0;
  args.debug = g_epilogue_debug;
  args.M = M; args.N = N; args.K = K;
  args.ldc = ldc; args.ldmask = ldmask;
  args.C = C; args.bias = bias;
  args.mask = reinterpret_cast<const __nv_bfloat16*>(relu_mask);
  args.flags = flags;
  args.m_tiles = (M + BM - 1) / BM;
  args.n_tiles = (N + bn - 1) / bn;
  args.k_blocks = (K + BK - 1) / BK;
  // Split-K: partial products of K slices are summed by the TMA reduce-add epilogue, so it needs
  // an fp32 output in ACCUM mode (C holds zeros or a running sum).  0 = pick: enough slices to
  // give every SM a work item, at least four k-blocks each.
  const bool can_split = tma_store && out_f32 && (flags & CB2_GEMM_ACCUM) && bias == nullptr &&
                         !(flags & CB2_GEMM_RELU);
  int splits = split_k;
  if (splits == 0) {
    splits = 1;
    const int tiles = args.m_tiles * args.n_tiles;
    if (can_split && tiles * 2 <= num_sms()) {
      splits = num_sms() / tiles;                 // one wave of work items
      if (splits > args.k_blocks / 2) splits = args.k_blocks / 2;
      if (splits < 1) splits = 1;
    }
  }
  if (splits > 1 && !can_split) {
    set_last_error("gemm: split_k=%d needs OUT_F32|ACCUM, a TMA-storable C and no bias/ReLU", splits);
    return CB2_ERR_ARG;
  }
  if (splits > args.k_blocks) splits = args.k_blocks;
  args.kb_per_split = (args.k_blocks + splits - 1) / splits;
  args.splits = (args.k_blocks + args.kb_per_split - 1) / args.kb_per_split;   // no empty slices
  args.colstats = colstats;
  args.m_fast = colstats != nullptr ? 1 : 0;
  args.conv = 0; args.conv_oh = 0; args.conv_a_bytes = 0;
  args.fin_out = reinterpret_cast<__nv_bfloat16*>(g_finalize.out);
  args.fin_ld = g_finalize.ld;
  args.fin_counters = g_finalize.counters;
  if (args.fin_out != nullptr) {
    if (!can_split) {
      set_last_error("gemm_splitk: workspace must be TMA-storable fp32");
      return CB2_ERR_ARG;
    }
    if (args.m_tiles * args.n_tiles * 4 > g_finalize.n_counters) {
      set_last_error("gemm_splitk: %d counters needed, %d given", args.m_tiles * args.n_tiles * 4,
                     g_finalize.n_counters);
      return CB2_ERR_ARG;
    }
  }
  if (colstats != nullptr && (!tma_store || out_f32 || bias != nullptr)) {
    set_last_error("gemm: column statistics need a bf16, TMA-storable C and no bias");
    return CB2_ERR_ARG;
  }
  // K-major, 128B swizzle: 8-row groups are 1024 B apart (SBO); LBO unused (=1);
  //                        one UMMA_K step = 16 elements = 32 B inside the swizzle span.
  // MN-major, 128B swizzle: 64-element MN chunks are BK*128 B apart (LBO); 8-k-row groups are
  //                        1024 B apart (SBO); one UMMA_K step = 16 k-rows = 2048 B.
  args.a_lbo = a_mn ? (BK * 128) >> 4 : 1;
  args.a_sbo = 1024 >> 4;
  args.a_kadv = a_mn ? (UMMA_K * 128) >> 4 : (UMMA_K * 2) >> 4;
  args.b_lbo = b_mn ? (BK * 128) >> 4 : 1;
  args.b_sbo = 1024 >> 4;
  args.b_kadv = b_mn ? (UMMA_K * 128) >> 4 : (UMMA_K * 2) >> 4;
  if (g_desc_override.on) {
    args.a_lbo = g_desc_override.a_lbo; args.a_sbo = g_desc_override.a_sbo;
    args.b_lbo = g_desc_override.b_lbo; args.b_sbo = g_desc_override.b_sbo;
    args.a_kadv = g_desc_override.a_kadv; args.b_kadv = g_desc_override.b_kadv;
  }

  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
#define CB2_DISPATCH(BNV)                                                            \
  if (!a_mn && !b_mn) return launch_gemm<BNV, false, false>(tmA, tmB, tmC, args, s); \
  if (!a_mn && b_mn) return launch_gemm<BNV, false, true>(tmA, tmB, tmC, args, s);   \
  if (a_mn && b_mn) return launch_gemm<BNV, true, true>(tmA, tmB, tmC, args, s);     \
  return launch_gemm<BNV, true, false>(tmA, tmB, tmC, args, s);
  if (bn == 64) { CB2_DISPATCH(64) }
  if (bn == 128) { CB2_DISPATCH(128) }
  CB2_DISPATCH(256)
#undef CB2_DISPATCH
}

// ---------------------------------------------------------------------------------------------
// ResNet stem: 7x7 / stride 2 / pad 3 convolution of a 3-channel image as an implicit GEMM.
//
// cuDNN needs 1.6 ms (forward) + 0.9 ms (weight gradient) for it at batch 256 - C = 3 defeats its
// tensor-core paths - against ~0.1 ms of memory traffic.  Here the image is first rearranged
// ("space to depth") into xs[B, OH+3, OW+3, 16]: pixel (r, s) of xs holds the 2x2 block of input
// pixels (2r-4+a, 2s-4+b), a, b in {0, 1}, as channels (a*2+b)*3 + c (12 used, 4 zero).  The 7x7/2
// convolution then is a 4x4 / stride-1 convolution of xs, and - the point of the layout - the 4
// pixels x 16 channels one kernel row touches are 128 contiguous bytes: exactly one row of a
// 128B-swizzled UMMA operand tile.  So the A operand needs no im2col buffer: a 4-D tensor map with
// a 32-byte stride between consecutive output pixels (overlapping rows) lets one TMA load fetch
// the 112 windows of one output row for one kernel row.
//   forward : Y[(b, oh), ow, :] = sum_i  A_i[(b, oh)] (112 x 64)  @  Wm[:, i*64 : i*64+64]^T
//   wgrad   : dWm^T[256, 64] = sum over (b, oh, ow)  window(b, oh, ow)^T  dY[b, oh, ow, :]
// Wm[64, 256] is the kernel in the same (i, j, a, b, c) order (cloud_b200/ops/conv.py builds it).
// ---------------------------------------------------------------------------------------------
static void fill_descriptors(GemmArgs& args, bool a_mn, bool b_mn) {
  args.a_lbo = a_mn ? (BK * 128) >> 4 : 1;
  args.a_sbo = 1024 >> 4;
  args.a_kadv = a_mn ? (UMMA_K * 128) >> 4 : (UMMA_K * 2) >> 4;
  args.b_lbo = b_mn ? (BK * 128) >> 4 : 1;
  args.b_sbo = 1024 >> 4;
  args.b_kadv = b_mn ? (UMMA_K * 128) >> 4 : (UMMA_K * 2) >> 4;
}

static int stem_check(const void* xs, int B, int OH, int OW, const char* what) {
  if (B <= 0 || OH <= 0 || OW <= 0 || OW > 128 || (reinterpret_cast<uintptr_t>(xs) & 15)) {
    set_last_error("%s: unsupported geometry B=%d OH=%d OW=%d (OW <= 128)", what, B, OH, OW);
    return CB2_ERR_UNSUPPORTED;
  }
  return 0;
}

static int stem_input_map(CUtensorMap* tm, const void* xs, int B, int OH, int OW, int box_ow) {
  const cuuint64_t px = 32;                                     // bytes per space-to-depth pixel
  const cuuint64_t dims[4] = {64, static_cast<cuuint64_t>(OW), static_cast<cuuint64_t>(OH + 3),
                              static_cast<cuuint64_t>(B)};
  const cuuint64_t strides[3] = {px, (OW + 3) * px, static_cast<cuuint64_t>(OH + 3) * (OW + 3) * px};
  const cuuint32_t box[4] = {64, static_cast<cuuint32_t>(box_ow), 1, 1};
  return make_tmap_nd(tm, xs, 4, dims, strides, box);
}

static int stem_output_map(CUtensorMap* tm, const void* y, int B, int OH, int OW, int box_ow) {
  const cuuint64_t dims[3] = {64, static_cast<cuuint64_t>(OW), static_cast<cuuint64_t>(B) * OH};
  const cuuint64_t strides[2] = {128, static_cast<cuuint64_t>(OW) * 128};
  const cuuint32_t box[3] = {64, static_cast<cuuint32_t>(box_ow), 1};
  return make_tmap_nd(tm, y, 3, dims, strides, box);
}

extern "C" int cb2_conv_stem_forward(const void* xs, const void* wm, void* y, int B, int OH, int OW,
                                     void* stream) {
  int rc = stem_check(xs, B, OH, OW, "conv_stem_forward");
  if (rc) return rc;
  CUtensorMap tmA, tmB, tmC;
  if ((rc = stem_input_map(&tmA, xs, B, OH, OW, OW))) return rc;
  if ((rc = make_tmap(&tmB, wm, 256, 64, 256, BK, 64))) return rc;
  if ((rc = stem_output_map(&tmC, y, B, OH, OW, 32))) return rc;
  GemmArgs args = {};
  args.M = B * OH * BM; args.N = 64; args.K = 256;
  args.ldc = 64; args.C = y;
  args.flags = 0;
  args.m_tiles = B * OH; args.n_tiles = 1; args.k_blocks = 4;
  args.splits = 1; args.kb_per_split = 4;
  args.tma_store = 1;
  // 4 = input rows kept in a ring of smem stages (each row fetched once per CTA instead of once per
  // kernel row) and the weights resident; TFC_STEM_RING=0 selects the plain per-tile loads (1)
  static const bool ring = [] { const char* e = getenv("TFC_STEM_RING"); return !(e && e[0] == '0'); }();
  args.conv = ring ? 4 : 1; args.conv_oh = OH; args.conv_a_bytes = OW * 128;
  fill_descriptors(args, false, false);
  return launch_gemm<64, false, false>(tmA, tmB, tmC, args, reinterpret_cast<cudaStream_t>(stream));
}

extern "C" int cb2_conv_stem_wgrad(const void* xs, const void* dy, float* dwt /*[256, 64] += */, int B,
                                   int OH, int OW, void* stream) {
  int rc = stem_check(xs, B, OH, OW, "conv_stem_wgrad");
  if (rc) return rc;
  CUtensorMap tmA, tmB, tmC;
  if ((rc = stem_input_map(&tmA, xs, B, OH, OW, 64))) return rc;
  if ((rc = stem_output_map(&tmB, dy, B, OH, OW, 64))) return rc;
  if ((rc = make_tmap(&tmC, dwt, 64, 256, 64, 32, 32, true))) return rc;
  GemmArgs args = {};
  args.M = 256; args.N = 64; args.K = B * OH * 128;
  args.ldc = 64; args.C = dwt;
  args.flags = CB2_GEMM_OUT_F32 | CB2_GEMM_ACCUM;
  args.m_tiles = 2; args.n_tiles = 1; args.k_blocks = B * OH * 2;
  int splits = num_sms() / 2;
  if (splits > args.k_blocks) splits = args.k_blocks;
  args.kb_per_split = (args.k_blocks + splits - 1) / splits;
  args.splits = (args.k_blocks + args.kb_per_split - 1) / args.kb_per_split;
  args.tma_store = 1;
  args.conv = 2; args.conv_oh = OH; args.conv_a_bytes = 0;
  fill_descriptors(args, true, true);
  return launch_gemm<64, true, true>(tmA, tmB, tmC, args, reinterpret_cast<cudaStream_t>(stream));
}

// ---------------------------------------------------------------------------------------------
// Weight gradient of a k x k, stride-1 convolution as split-K GEMMs over shifted pixel boxes.
//   dW[(tap, ci), co] = sum over output pixels (b, h, w) of x[b, h + i - pad, w + j - pad, ci] * dy[b, h, w, co]
// Both operands are read in place from the NHWC tensors through 4-D tensor maps {C, W, H, B}: the
// GEMM-K block is a box of 64 output pixels (w_box x h_box, w_box = next power of two >= W), the A
// chunk for tap (i, j) is the same box moved by (i - pad, j - pad), and every pixel that falls
// outside the image - the zero padding of the convolution, the box tail beyond W or H - comes back
// as zeros from TMA.  No im2col buffer, no padded copy of the activation.
// Workload: the sixteen 3x3 convolutions of ResNet-50 (cuDNN falls back to an sm_80 kernel for the
// 56x56 ones: 0.77 ms each at batch 256).
// ---------------------------------------------------------------------------------------------
extern "C" int cb2_conv_wgrad(const void* x, const void* dy, float* dwt /*[taps*Cin (+pad), Cout] += */,
                              int B, int H, int W, int Cin, int Cout, int ksize, int pad, int stride,
                              void* stream) {
  // H, W: OUTPUT size; the input is [(H-1)*stride + 1 .. H*stride] rows - passed as Hin = H*stride
  // for even inputs (the only case the front end produces): pixels (h*stride + i - pad, ...)
  const int Hin = H * stride, Win = W * stride;
  if (B <= 0 || H <= 0 || W <= 0 || W > 64 || (Cin & 63) || (Cout & 63) || ksize <= 0 || ksize > 7 ||
      pad != ksize / 2 || stride < 1 || stride > 2 || (reinterpret_cast<uintptr_t>(x) & 15) ||
      (reinterpret_cast<uintptr_t>(dy) & 15)) {
    set_last_error("conv_wgrad: unsupported geometry B=%d H=%d W=%d Cin=%d Cout=%d k=%d pad=%d stride=%d "
                   "(output W <= 64, channels multiples of 64, pad = k/2, stride 1 or 2)",
                   B, H, W, Cin, Cout, ksize, pad, stride);
    return CB2_ERR_UNSUPPORTED;
  }
  int w_box = 8;
  while (w_box < W) w_box <<= 1;
  const int h_box = 64 / w_box;
  const int hblocks = (H + h_box - 1) / h_box;
  const int taps = ksize * ksize;
  const int bn = Cout >= 256 ? 256 : (Cout >= 128 ? 128 : 64);
  auto pixel_map = [&](CUtensorMap* tm, const void* base, int C, int hh, int ww, int estride) {
    const cuuint64_t dims[4] = {static_cast<cuuint64_t>(C), static_cast<cuuint64_t>(ww),
                                static_cast<cuuint64_t>(hh), static_cast<cuuint64_t>(B)};
    const cuuint64_t strides[3] = {static_cast<cuuint64_t>(C) * 2, static_cast<cuuint64_t>(ww) * C * 2,
                                   static_cast<cuuint64_t>(hh) * ww * C * 2};
    // with a traversal stride the box spans w_box*stride pixels and every stride-th one is loaded
    const cuuint32_t box[4] = {64, static_cast<cuuint32_t>(w_box * estride),
                               static_cast<cuuint32_t>(h_box * estride), 1};
    const cuuint32_t estr[4] = {1, static_cast<cuuint32_t>(estride), static_cast<cuuint32_t>(estride), 1};
    EncodeTiledFn fn = get_encode_fn();
    if (!fn) {
      set_last_error("cuTensorMapEncodeTiled entry point not available");
      return static_cast<int>(CB2_ERR_CUDA);
    }
    CUresult r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                    CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
      set_last_error("conv_wgrad: cuTensorMapEncodeTiled failed (%d)", static_cast<int>(r));
      return static_cast<int>(CB2_ERR_CUDA);
    }
    return 0;
  };
  CUtensorMap tmA, tmB, tmC;
  int rc;
  if ((rc = pixel_map(&tmA, x, Cin, Hin, Win, stride))) return rc;
  if ((rc = pixel_map(&tmB, dy, Cout, H, W, 1))) return rc;
  const int rows = taps * Cin;
  if ((rc = make_tmap(&tmC, dwt, Cout, rows, Cout, 32, 32, true))) return rc;
  GemmArgs args = {};
  args.M = rows; args.N = Cout; args.K = B * hblocks * 64;
  args.ldc = Cout; args.C = dwt;
  args.flags = CB2_GEMM_OUT_F32 | CB2_GEMM_ACCUM;
  args.m_tiles = (rows + BM - 1) / BM; args.n_tiles = (Cout + bn - 1) / bn;
  args.k_blocks = B * hblocks;
  const int tiles = args.m_tiles * args.n_tiles;
  int splits = num_sms() / tiles;                     // one wave of work items
  if (splits > args.k_blocks / 2) splits = args.k_blocks / 2;
  if (splits < 1) splits = 1;
  args.kb_per_split = (args.k_blocks + splits - 1) / splits;
  args.splits = (args.k_blocks + args.kb_per_split - 1) / args.kb_per_split;
  args.tma_store = 1;
  args.conv = 3;
  args.conv_cin = Cin; args.conv_taps = taps; args.conv_kw = ksize; args.conv_pad = pad;
  args.conv_hblocks = hblocks; args.conv_hbox = h_box; args.conv_stride = stride;
  fill_descriptors(args, true, true);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (bn == 64) return launch_gemm<64, true, true>(tmA, tmB, tmC, args, st);
  if (bn == 128) return launch_gemm<128, true, true>(tmA, tmB, tmC, args, st);
  return launch_gemm<256, true, true>(tmA, tmB, tmC, args, st);
}
