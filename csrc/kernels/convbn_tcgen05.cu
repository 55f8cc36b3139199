// 1x1 convolution + batch normalisation folded into GEMMs (B200, sm_100a).
//
// ResNet's wide 1x1 convolutions (K = Cin small, N = Cout = 4 Cin) are bound by the bytes of their
// output, and the batch-norm behind them reads and writes that output several more times.  Here the
// convolution output y = X W^T is never stored:
//
//   forward   statistics of y follow from the Gram matrix of the (small) input:
//             mean_c = w_c . colsum(X) / M,  E[y_c^2] = w_c^T (X^T X) w_c / M   (cb2_convbn_fwd_finalize)
//             then ONE GEMM whose epilogue applies scale/shift (+ shortcut) (+ ReLU) while the tile
//             is still in TMEM and writes the block output z (+ one sign bit per element),
//   backward  with g = dL/dz masked by the ReLU bits, P = g^T X (a split-K GEMM that exists anyway)
//             gives sum(g xhat) = rstd (sum_k P W - mean sum(g)), so the batch-norm reductions
//             never touch y, and
//               dW = s'(P - a S - b rstd (W Gram - mean S)),
//               dX = g (s' W) - X (W^T diag(t) W) - const                (cb2_convbn_bwd_finalize)
//             i.e. the input gradient is ONE GEMM over two K sources ([g | X] against [s' W ; -Q]).
//
// This file holds the GEMM those steps share: a persistent, warp-specialised tcgen05 kernel
// (TMA -> 128B-swizzled smem -> tcgen05.mma -> TMEM, double-buffered accumulators) with
//   * up to two (A, B) operand pairs walked back to back along K,
//   * eight epilogue warps (two per TMEM lane quarter, each owning half of the tile's columns),
//   * an epilogue that computes  out = mask_in ? relu?(acc * scale + shift + residual) : 0,
//     the residual tile arriving through its own TMA ring filled by a dedicated warp,
//   * optional sign bits out, optional column sums of the stored values (the next layer's sum(g)).
//
// No counterpart in the reference (TensorFlow runs cuDNN convolution, batch-norm, add and ReLU
// kernels for core/tests/examples/call_run_within_script_with_keras_fit.py:79-90).
#include <stdlib.h>

#include "gemm_host.cuh"

namespace cb2 {

constexpr int FBM = 128;
constexpr int FBK = 64;
constexpr int kFusedThreads = 384;          // warps: 0 TMA, 1 MMA, 2 residual TMA, 3 idle, 4..11 epilogue
constexpr int kFAStage = FBM * FBK * 2;     // 16 KiB
constexpr int kResSlot = FBM * 128;         // 128 rows x 64 bf16
constexpr int kFStatCopies = 8;             // = kColStatCopies / kBnCopies
constexpr int kMaxStages = 8;
constexpr int kMaxRing = 8;

struct FusedArgs {
  int M, N;
  int kb1, kb2;                 // 64-wide k blocks of operand pair 1 / pair 2
  int m_tiles, n_tiles;
  int stages, ring;             // smem pipeline depth; residual ring slots (0: no residual)
  int out_bufs;                 // staging buffers per epilogue warp (2, or 1 when the ring needs the room)
  uint32_t b_lbo, b_sbo, b_kadv;
  const float* scale;           // nullable [N]
  const float* shift;           // nullable [N]
  int relu;
  const uint32_t* mask_in;      // nullable, bits [M, N/32]: element kept where its bit is set
  uint32_t* mask_out;           // nullable, bits [M, N/32]: stored value > 0
  float* colstats;              // nullable [kFStatCopies][2][N]
};

struct alignas(64) FusedMaps {
  CUtensorMap a1, b1, a2, b2, c, r;
};

template <int BN>
struct FusedCfg {
  static constexpr int kEW = BN >= 128 ? 8 : 4;          // active epilogue warps
  static constexpr int kHF = kEW / 4;                    // column halves
  static constexpr int kNC = BN / 64;                    // 64-column chunks per tile
  static constexpr int kNCW = kNC / kHF;                 // chunks per epilogue warp
  static constexpr int kBStage = BN * FBK * 2;
  static constexpr int kStage = kFAStage + kBStage;
  static constexpr int staging(int out_bufs) { return kEW * out_bufs * 4096; }
  static constexpr int kParams = 8 * 2 * 128 * 4;        // per-warp scale / shift copies
  static constexpr int kBars = 512;
  static constexpr int kTmemCols = 2 * BN;
  static constexpr int smem_bytes(int stages, int ring, int out_bufs = 2) {
    return stages * kStage + staging(out_bufs) + ring * kResSlot + kParams + kBars + 1024;
  }
};

template <int BN, bool B_MN>
__global__ void __launch_bounds__(kFusedThreads, 1)
convbn_gemm_kernel(const __grid_constant__ FusedMaps tm, const FusedArgs p) {
  using Cfg = FusedCfg<BN>;
  constexpr int kEW = Cfg::kEW, kHF = Cfg::kHF, kNC = Cfg::kNC, kNCW = Cfg::kNCW;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  const int S = p.stages, R = p.ring;
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem_a + S * kFAStage;
  uint8_t* smem_out = smem_b + S * Cfg::kBStage;
  uint8_t* smem_res = smem_out + Cfg::staging(p.out_bufs);
  uint8_t* smem_par = smem_res + R * kResSlot;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_par + Cfg::kParams);
  uint64_t* full_bar = bars;                               // [kMaxStages]
  uint64_t* empty_bar = bars + kMaxStages;                 // [kMaxStages]
  uint64_t* tfull_bar = bars + 2 * kMaxStages;             // [2]
  uint64_t* tempty_bar = bars + 2 * kMaxStages + 2;        // [2]
  uint64_t* rfull_bar = bars + 2 * kMaxStages + 4;         // [kMaxRing][2]: fills of a slot alternate
  uint64_t* rempty_bar = bars + 2 * kMaxStages + 4 + 2 * kMaxRing;   // [kMaxRing]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * kMaxStages + 4 + 3 * kMaxRing);

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);    // warp-uniform for ptxas
  const int lane = threadIdx.x & 31;
  const bool has_res = R > 0;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tm.a1);
    tma_prefetch_desc(&tm.b1);
    tma_prefetch_desc(&tm.c);
    if (p.kb2 > 0) {
      tma_prefetch_desc(&tm.a2);
      tma_prefetch_desc(&tm.b2);
    }
    if (has_res) tma_prefetch_desc(&tm.r);
    for (int s = 0; s < kMaxStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tfull_bar[a], 1);
      mbar_init(&tempty_bar[a], kEW);
    }
    for (int s = 0; s < kMaxRing; ++s) {
      mbar_init(&rfull_bar[2 * s], 1);
      mbar_init(&rfull_bar[2 * s + 1], 1);
      mbar_init(&rempty_bar[s], 4);
    }
    mbar_fence_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, Cfg::kTmemCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const int num_tiles = p.m_tiles * p.n_tiles;
  const int kbt = p.kb1 + p.kb2;
  // m tile index fastest: a CTA keeps the same column block (scale / shift, statistics) for long runs

  if (warp == 0) {
    // ===================== TMA producer: operands =====================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const int m0 = (tile % p.m_tiles) * FBM, n0 = (tile / p.m_tiles) * BN;
        for (int kb = 0; kb < kbt; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem_a + stage * kFAStage;
          uint8_t* sb = smem_b + stage * Cfg::kBStage;
          const bool second = kb >= p.kb1;
          const CUtensorMap* ta = second ? &tm.a2 : &tm.a1;
          const CUtensorMap* tb = second ? &tm.b2 : &tm.b1;
          const int k0 = (second ? kb - p.kb1 : kb) * FBK;
          mbar_expect_tx(&full_bar[stage], Cfg::kStage);
          tma_load_2d(ta, &full_bar[stage], sa, k0, m0);                 // box {64 k, 128 rows}
          if (!B_MN) {
            tma_load_2d(tb, &full_bar[stage], sb, k0, n0);               // box {64 k, BN rows}
          } else {
#pragma unroll
            for (int c = 0; c < BN / 64; ++c)                            // box {64 n, 64 k}
              tma_load_2d(tb, &full_bar[stage], sb + c * (FBK * 128), n0 + c * 64, k0);
          }
          if (++stage == S) { stage = 0; phase ^= 1; }
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ===================== MMA issuer (whole warp, one elected lane issues: common.cuh) ==========
    {
      const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((B_MN ? 1u : 0u) << 16) |
                             (static_cast<uint32_t>(BN >> 3) << 17) |
                             (static_cast<uint32_t>(FBM >> 4) << 24);
      int stage = 0;
      uint32_t phase = 0;
      int t = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++t) {
        const int acc = t & 1;
        mbar_wait_warp(&tempty_bar[acc], ((t >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + acc * BN;
        for (int kb = 0; kb < kbt; ++kb) {
          mbar_wait_warp(&full_bar[stage], phase);
          tc_fence_after();
          const uint64_t adesc = make_smem_desc(smem_u32(smem_a + stage * kFAStage), 1, 1024 >> 4);
          const uint64_t bdesc =
              make_smem_desc(smem_u32(smem_b + stage * Cfg::kBStage), p.b_lbo, p.b_sbo);
#pragma unroll
          for (int k = 0; k < FBK / 16; ++k)
            umma_bf16_warp(tmem_d, adesc + static_cast<uint64_t>(k * 2),
                           bdesc + static_cast<uint64_t>(k * p.b_kadv), idesc, (kb | k) != 0);
          umma_commit_warp(&empty_bar[stage]);
          if (kb == kbt - 1) umma_commit_warp(&tfull_bar[acc]);
          if (++stage == S) { stage = 0; phase ^= 1; }
        }
      }
    }
    __syncwarp();
  } else if (warp == 2) {
    // ===================== TMA producer: residual tiles =====================
    if (lane == 0 && has_res) {
      int seq = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const int m0 = (tile % p.m_tiles) * FBM, n0 = (tile / p.m_tiles) * BN;
        // the order in which the epilogue warps ask for them: chunk j of column half 0, of half 1, ...
        for (int j = 0; j < kNCW; ++j)
          for (int hf = 0; hf < kHF; ++hf, ++seq) {
            const int c0 = n0 + (hf * kNCW + j) * 64;
            const int slot = seq % R;
            const int use = seq / R;                         // n-th fill of this slot
            uint64_t* rfull = &rfull_bar[2 * slot + (use & 1)];
            mbar_wait(&rempty_bar[slot], (use & 1) ^ 1);
            if (c0 < p.N) {
              mbar_expect_tx(rfull, kResSlot);
              tma_load_2d(&tm.r, rfull, smem_res + slot * kResSlot, c0, m0);
            } else {
              mbar_arrive(rfull);                  // chunk past the last column: nothing to fetch
            }
          }
      }
    }
    __syncwarp();
  } else if (warp >= 4 && warp < 4 + kEW) {
    // ===================== epilogue =====================
    const int q = warp & 3;
    const int hf = (warp - 4) >> 2;
    const uint32_t sw = lane & 7;
    const uint32_t stg = smem_u32(smem_out) + (warp - 4) * p.out_bufs * 4096;
    const uint32_t par = smem_u32(smem_par) + (warp - 4) * 1024;     // [0,512): scale, [512,1024): shift
    const bool has_scale = p.scale != nullptr, has_shift = p.shift != nullptr;
    const bool affine = has_scale || has_shift;
    const bool relu = p.relu != 0;
    const bool has_min = p.mask_in != nullptr, has_mout = p.mask_out != nullptr;
    const int words_per_row = p.N >> 5;
    int stage_buf = 0;
    int par_n0 = -1;
    float cs[kNCW][4];
    int cs_n0 = -1;
#pragma unroll
    for (int j = 0; j < kNCW; ++j) cs[j][0] = cs[j][1] = cs[j][2] = cs[j][3] = 0.f;
    int t = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++t) {
      const int acc = t & 1;
      const int m0 = (tile % p.m_tiles) * FBM, n0 = (tile / p.m_tiles) * BN;
      const int row = m0 + q * 32 + lane;
      const int cw0 = n0 + hf * kNCW * 64;                   // first column this warp owns
      if (affine && par_n0 != n0) {
        par_n0 = n0;
        __syncwarp();
        for (int i = lane; i < kNCW * 64; i += 32) {
          const int c = cw0 + i;
          const float scv = (has_scale && c < p.N) ? __ldg(p.scale + c) : 1.f;
          const float shv = (has_shift && c < p.N) ? __ldg(p.shift + c) : 0.f;
          asm volatile("st.shared.f32 [%0], %1;" ::"r"(par + i * 4), "f"(scv) : "memory");
          asm volatile("st.shared.f32 [%0], %1;" ::"r"(par + 512 + i * 4), "f"(shv) : "memory");
        }
        __syncwarp();
      }
      if (p.colstats != nullptr && cs_n0 != n0) {
        if (cs_n0 >= 0) {
          float* dst = p.colstats + static_cast<size_t>((blockIdx.x * 4 + q) % kFStatCopies) * 2 * p.N;
#pragma unroll
          for (int j = 0; j < kNCW; ++j) {
            const int n = cs_n0 + (hf * kNCW + j) * 64 + 2 * lane;
            if (n < p.N) {
              atomicAdd(dst + n, cs[j][0]);
              atomicAdd(dst + n + 1, cs[j][1]);
              atomicAdd(dst + p.N + n, cs[j][2]);
              atomicAdd(dst + p.N + n + 1, cs[j][3]);
            }
            cs[j][0] = cs[j][1] = cs[j][2] = cs[j][3] = 0.f;
          }
        }
        cs_n0 = n0;
      }
      // ReLU bits of this row for the columns this warp owns (2 words per chunk)
      uint32_t min_w[kNCW * 2], mout_w[kNCW * 2];
#pragma unroll
      for (int i = 0; i < kNCW * 2; ++i) { min_w[i] = 0xffffffffu; mout_w[i] = 0u; }
      if (p.mask_in != nullptr && row < p.M) {
        const uint32_t* mp = p.mask_in + static_cast<size_t>(row) * words_per_row + (cw0 >> 5);
#pragma unroll
        for (int i = 0; i < kNCW * 2; ++i)
          if (cw0 + i * 32 < p.N) min_w[i] = __ldg(mp + i);
      }

      mbar_wait(&tfull_bar[acc], (t >> 1) & 1);
      tc_fence_after();
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * BN + hf * kNCW * 64;

#pragma unroll
      for (int j = 0; j < kNCW; ++j) {
        const int c0 = cw0 + j * 64;
        // Position in the residual ring order.  The ring is shared by the two column halves, and with
        // an odd length the fills of one slot alternate between them - so a warp sees only every
        // other fill of a slot.  A parity wait is only sound for a waiter that sees EVERY phase of its
        // barrier (else "two phases ago" reads as "done": the warp would take a slot whose previous
        // fill, a tile of the other half, is still in flight, and release it early).  Hence two "full"
        // barriers per slot, used by alternate fills: every waiter of a barrier sees all its phases.
        const int seq = (t * kNCW + j) * kHF + hf;
        const int slot = has_res ? seq % R : 0;
        const int use = has_res ? seq / R : 0;
        uint64_t* rfull = &rfull_bar[2 * slot + (use & 1)];
        const uint32_t ring_phase = (use >> 1) & 1;
        if (c0 < p.N) {
          uint32_t v[2][32];
          tmem_ld_32x32(taddr + j * 64, v[0]);
          tmem_ld_32x32(taddr + j * 64 + 32, v[1]);
          if (has_res) mbar_wait(rfull, ring_phase);
          const uint32_t buf = stg + stage_buf * 4096;
          const uint32_t rowp = buf + lane * 128;
          const uint32_t resp = smem_u32(smem_res) + slot * kResSlot + (q * 32 + lane) * 128;
          if (lane == 0) {                                  // the store that last read this buffer is done
            if (p.out_bufs == 2) tma_store_wait_read<1>();
            else tma_store_wait_read<0>();
          }
          __syncwarp();
          tmem_ld_wait_for(v[0]);
          reg_fence(v[1]);
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            const uint32_t mi = min_w[j * 2 + h];
            uint32_t mo = 0;
#pragma unroll
            for (int g = 0; g < 4; ++g) {
              float f[8];
#pragma unroll
              for (int e = 0; e < 8; ++e) f[e] = __uint_as_float(v[h][g * 8 + e]);
              if (affine) {
                const uint32_t po = (j * 64 + h * 32 + g * 8) * 4;
                const float4 s0 = ld_shared_f4(par + po), s1 = ld_shared_f4(par + po + 16);
                const float4 h0 = ld_shared_f4(par + 512 + po), h1 = ld_shared_f4(par + 512 + po + 16);
                f[0] = fmaf(f[0], s0.x, h0.x); f[1] = fmaf(f[1], s0.y, h0.y);
                f[2] = fmaf(f[2], s0.z, h0.z); f[3] = fmaf(f[3], s0.w, h0.w);
                f[4] = fmaf(f[4], s1.x, h1.x); f[5] = fmaf(f[5], s1.y, h1.y);
                f[6] = fmaf(f[6], s1.z, h1.z); f[7] = fmaf(f[7], s1.w, h1.w);
              }
              if (has_res) {
                uint32_t r0, r1, r2, r3;
                asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];"
                             : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
                             : "r"(resp + (((h * 4 + g) ^ sw) << 4)));
                f[0] += bf16_lo(r0); f[1] += bf16_hi(r0); f[2] += bf16_lo(r1); f[3] += bf16_hi(r1);
                f[4] += bf16_lo(r2); f[5] += bf16_hi(r2); f[6] += bf16_lo(r3); f[7] += bf16_hi(r3);
              }
              if (has_min) {
                const uint32_t mb = mi >> (g * 8);
#pragma unroll
                for (int e = 0; e < 8; ++e)
                  if (!((mb >> e) & 1u)) f[e] = 0.f;
              }
              if (relu) {
#pragma unroll
                for (int e = 0; e < 8; ++e) f[e] = fmaxf(f[e], 0.f);
              }
              if (has_mout) {
#pragma unroll
                for (int e = 0; e < 8; ++e) mo |= (f[e] > 0.f ? 1u : 0u) << (g * 8 + e);
              }
              st_shared_v4(rowp + (((h * 4 + g) ^ sw) << 4), pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]),
                           pack_bf16(f[4], f[5]), pack_bf16(f[6], f[7]));
            }
            mout_w[j * 2 + h] = mo;
          }
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) {
            tma_store_2d(&tm.c, buf, c0, m0 + q * 32);
            tma_store_commit();
            if (has_res) mbar_arrive(&rempty_bar[slot]);     // all 32 lanes have read their residual row
          }
          if (p.colstats != nullptr) {
            // column sums of the bf16 values just staged: lane l owns columns 2l, 2l+1 of the chunk
            const uint32_t colp = buf + ((lane & 3) << 2);
            const uint32_t jj = lane >> 2;
#pragma unroll
            for (int r0 = 0; r0 < 32; r0 += 16) {
              uint32_t wv[16];
#pragma unroll
              for (int r = 0; r < 16; ++r)
                asm volatile("ld.shared.b32 %0, [%1];"
                             : "=r"(wv[r])
                             : "r"(colp + (r0 + r) * 128 + ((jj ^ ((r0 + r) & 7)) << 4)));
#pragma unroll
              for (int r = 0; r < 16; ++r) {
                // rows past M hold TMEM garbage * 0: they were computed from zero-filled A rows,
                // i.e. shift (+ 0 residual): exclude them explicitly
                const bool ok = m0 + q * 32 + r0 + r < p.M;
                const float a = ok ? __uint_as_float(wv[r] << 16) : 0.f;
                const float b = ok ? __uint_as_float(wv[r] & 0xffff0000u) : 0.f;
                cs[j][0] += a;
                cs[j][1] += b;
                cs[j][2] = fmaf(a, a, cs[j][2]);
                cs[j][3] = fmaf(b, b, cs[j][3]);
              }
            }
          }
          stage_buf = (stage_buf + 1) & (p.out_bufs - 1);
        } else if (has_res) {
          // keep the ring in step for a chunk past the last column
          mbar_wait(rfull, ring_phase);
          __syncwarp();
          if (lane == 0) mbar_arrive(&rempty_bar[slot]);
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty_bar[acc]);
      if (p.mask_out != nullptr && row < p.M) {
        uint32_t* mp = p.mask_out + static_cast<size_t>(row) * words_per_row + (cw0 >> 5);
#pragma unroll
        for (int i = 0; i < kNCW * 2; ++i)
          if (cw0 + i * 32 < p.N) mp[i] = mout_w[i];
      }
    }
    if (p.colstats != nullptr && cs_n0 >= 0) {
      float* dst = p.colstats + static_cast<size_t>((blockIdx.x * 4 + q) % kFStatCopies) * 2 * p.N;
#pragma unroll
      for (int j = 0; j < kNCW; ++j) {
        const int n = cs_n0 + (hf * kNCW + j) * 64 + 2 * lane;
        if (n < p.N) {
          atomicAdd(dst + n, cs[j][0]);
          atomicAdd(dst + n + 1, cs[j][1]);
          atomicAdd(dst + p.N + n, cs[j][2]);
          atomicAdd(dst + p.N + n + 1, cs[j][3]);
        }
      }
    }
    if (lane == 0) tma_store_wait<0>();   // staging smem must outlive its stores
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, Cfg::kTmemCols);
}

template <int BN, bool B_MN>
static int launch_fused(const FusedMaps& maps, FusedArgs& args, cudaStream_t stream) {
  using Cfg = FusedCfg<BN>;
  auto kern = convbn_gemm_kernel<BN, B_MN>;
  static bool configured = false;
  if (!configured) {
    CB2_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448));
    configured = true;
  }
  const int kbt = args.kb1 + args.kb2;
  int stages = (232448 - Cfg::smem_bytes(0, args.ring)) / Cfg::kStage;
  if (stages > kMaxStages) stages = kMaxStages;
  if (stages > 2 * kbt) stages = 2 * kbt > 2 ? 2 * kbt : 2;       // two tiles' worth is enough
  if (stages < 2) {
    set_last_error("convbn_gemm: shared memory budget leaves fewer than 2 pipeline stages");
    return CB2_ERR_UNSUPPORTED;
  }
  args.stages = stages;
  args.out_bufs = 2;
  if (args.ring > 0) {
    // the shortcut tiles are the largest read of an output-bound GEMM: whatever shared memory the
    // operand pipeline leaves goes into their ring (bytes in flight = bandwidth x latency)
    static const int ring_env = [] { const char* e = getenv("TFC_CONVBN_RING"); return e ? atoi(e) : 0; }();
    static const int bufs_env = [] { const char* e = getenv("TFC_CONVBN_OUT_BUFS"); return e ? atoi(e) : 0; }();
    if (bufs_env == 1) args.out_bufs = 1;        // tuning knob: the second staging buffers go to the ring
    int ring = (232448 - Cfg::smem_bytes(stages, 0, args.out_bufs)) / kResSlot;
    if (ring > kMaxRing) ring = kMaxRing;
    if (ring_env >= 1 && ring_env <= ring) ring = ring_env;           // tuning knob (bench/convbn_probe.py)
    if (ring < 1) {
      set_last_error("convbn_gemm: no room for the shortcut ring");
      return CB2_ERR_UNSUPPORTED;
    }
    args.ring = ring;
  }
  const int tiles = args.m_tiles * args.n_tiles;
  const int grid = tiles < num_sms() ? tiles : num_sms();
  kern<<<grid, kFusedThreads, Cfg::smem_bytes(stages, args.ring, args.out_bufs), stream>>>(maps, args);
  return check_cuda(cudaGetLastError(), "convbn_gemm launch");
}

}  // namespace cb2

using namespace cb2;

extern "C" int cb2_convbn_stat_copies(void) { return kFStatCopies; }

extern "C" int cb2_convbn_gemm(const void* A, int lda, const void* B, int ldb, int b_mn, const void* A2,
                               int lda2, const void* B2, int ldb2, int K2, void* C, int ldc, int M, int N,
                               int K1, const float* scale, const float* shift, const void* residual,
                               int ldr, int relu, const void* mask_in, void* mask_out, float* colstats,
                               int block_n, void* stream) {
  if (M <= 0 || N <= 0 || K1 <= 0 || K2 < 0) {
    set_last_error("convbn_gemm: empty problem M=%d N=%d K1=%d K2=%d", M, N, K1, K2);
    return CB2_ERR_ARG;
  }
  if ((N & 63) || (K1 & 63) || (K2 & 63)) {
    set_last_error("convbn_gemm: N, K1 and K2 must be multiples of 64 (N=%d K1=%d K2=%d)", N, K1, K2);
    return CB2_ERR_UNSUPPORTED;
  }
  auto bad = [](const void* p, int ld) { return (reinterpret_cast<uintptr_t>(p) & 15) || (ld & 7); };
  if (bad(A, lda) || bad(B, ldb) || bad(C, ldc) || (K2 > 0 && (bad(A2, lda2) || bad(B2, ldb2))) ||
      (residual != nullptr && bad(residual, ldr))) {
    set_last_error("convbn_gemm: TMA needs 16-byte aligned bases and row pitches that are multiples of 8");
    return CB2_ERR_ARG;
  }
  if (K2 > 0 && (A2 == nullptr || B2 == nullptr)) {
    set_last_error("convbn_gemm: K2 > 0 needs a second operand pair");
    return CB2_ERR_ARG;
  }
  const int m_tiles = (M + FBM - 1) / FBM;
  int bn = block_n;
  static const int bn_env = [] { const char* e = getenv("TFC_CONVBN_BN"); return e ? atoi(e) : 0; }();
  if (bn == 0 && (bn_env == 64 || bn_env == 128 || bn_env == 256) && bn_env <= N) bn = bn_env;
  if (bn == 0) {
    // the widest tile (fewest operand re-reads, best MMA shape) that keeps the waves full
    const int sms = num_sms();
    bn = 64;
    for (int cand : {256, 128}) {
      if (cand > N) continue;
      const long tiles = static_cast<long>(m_tiles) * ((N + cand - 1) / cand);
      const long waves = (tiles + sms - 1) / sms;
      if (tiles * 100 >= waves * sms * 80) { bn = cand; break; }    // >= 80 % of the slots do work
    }
  }
  if (bn != 64 && bn != 128 && bn != 256) {
    set_last_error("convbn_gemm: block_n must be 64, 128 or 256 (got %d)", bn);
    return CB2_ERR_ARG;
  }
  FusedMaps maps;
  int rc;
  if ((rc = make_tmap(&maps.a1, A, K1, M, lda, FBK, FBM))) return rc;
  if (!b_mn) rc = make_tmap(&maps.b1, B, K1, N, ldb, FBK, bn);
  else       rc = make_tmap(&maps.b1, B, N, K1, ldb, 64, FBK);
  if (rc) return rc;
  maps.a2 = maps.a1;
  maps.b2 = maps.b1;
  if (K2 > 0) {
    if ((rc = make_tmap(&maps.a2, A2, K2, M, lda2, FBK, FBM))) return rc;
    if (!b_mn) rc = make_tmap(&maps.b2, B2, K2, N, ldb2, FBK, bn);
    else       rc = make_tmap(&maps.b2, B2, N, K2, ldb2, 64, FBK);
    if (rc) return rc;
  }
  if ((rc = make_tmap(&maps.c, C, N, M, ldc, 64, 32))) return rc;
  maps.r = maps.c;
  if (residual != nullptr && (rc = make_tmap(&maps.r, residual, N, M, ldr, 64, FBM))) return rc;

  FusedArgs args;
  args.M = M; args.N = N;
  args.kb1 = K1 / FBK; args.kb2 = K2 / FBK;
  args.m_tiles = m_tiles; args.n_tiles = (N + bn - 1) / bn;
  args.ring = residual != nullptr ? 3 : 0;
  args.stages = 0;
  args.b_lbo = b_mn ? (FBK * 128) >> 4 : 1;
  args.b_sbo = 1024 >> 4;
  args.b_kadv = b_mn ? (16 * 128) >> 4 : (16 * 2) >> 4;
  args.scale = scale; args.shift = shift; args.relu = relu;
  args.mask_in = reinterpret_cast<const uint32_t*>(mask_in);
  args.mask_out = reinterpret_cast<uint32_t*>(mask_out);
  args.colstats = colstats;
  if ((mask_in != nullptr && (reinterpret_cast<uintptr_t>(mask_in) & 3)) ||
      (mask_out != nullptr && (reinterpret_cast<uintptr_t>(mask_out) & 3))) {
    set_last_error("convbn_gemm: mask pointers must be 4-byte aligned");
    return CB2_ERR_ARG;
  }
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (bn == 64) return b_mn ? launch_fused<64, true>(maps, args, st) : launch_fused<64, false>(maps, args, st);
  if (bn == 128) return b_mn ? launch_fused<128, true>(maps, args, st) : launch_fused<128, false>(maps, args, st);
  return b_mn ? launch_fused<256, true>(maps, args, st) : launch_fused<256, false>(maps, args, st);
}
