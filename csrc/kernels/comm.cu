// Peer-memory collectives for NVLink 5 / NVSwitch, fused with the optimizer update.
//
// Every rank maps every other rank's gradient arena, parameter arena and signal pad (CUDA VMM /
// IPC, see csrc/runtime/symm_arena.cc).  One kernel per gradient bucket then does
//
//   barrier -> reduce (peer loads | multimem.ld_reduce in the switch) -> scale -> Adam/SGD/RMSprop
//           -> publish parameters (local store | peer stores | multimem.st) -> barrier -> zero grads
//
// so gradients never bounce through a staging buffer and the optimizer costs no extra launch or
// HBM pass.  Three schedules:
//   ONESHOT  every rank reads all W arenas and updates a replicated state (small buckets: one
//            NVLink round trip, bitwise identical results on all ranks since the sum order is fixed)
//   TWOSHOT  rank r reduces and updates shard r only (ZeRO-1 style sharded state) and stores the
//            new parameters into all W parameter arenas
//   NVLS     as TWOSHOT but the sum happens in the NVSwitch (multimem.ld_reduce) and the
//            parameter all-gather is one multimem.st per 16 bytes
//
// Flag protocol: per (block, source-rank) 32-bit epoch slots in each rank's signal pad, written
// with st.release.sys and polled with ld.acquire.sys; epochs only grow, so no reset and no ABA.
// Every spin is bounded by a watchdog (common.cuh) so a dead peer traps instead of hanging.
//
// Reference parity: K9/K10/K11 of SURVEY.md section 2.5 - the all-reduce TensorFlow runs for
// MirroredStrategy()/MultiWorkerMirroredStrategy() (core/preprocess.py:139-143) plus the
// per-variable ResourceApplyAdam launches that follow it.
#include <stdlib.h>

#include "optim.cuh"

namespace cb2 {

struct DevCtx {
  int rank, world;
  void* grad[CB2_MAX_RANKS];
  void* param[CB2_MAX_RANKS];
  uint32_t* pad[CB2_MAX_RANKS];
  void* mc_grad;
  void* mc_param;
  uint32_t* epochs;
  uint4* ll[CB2_MAX_RANKS];      // every rank's LL receive buffer (flag-in-payload protocol)
  size_t ll_bytes;               // size of one such buffer
  uint32_t* ll_state;            // local: [0] calls completed, [1] blocks of the running call done
  unsigned long long timeout_ns; // how long a wait for a peer may last before the kernel traps
};

__device__ __forceinline__ unsigned long long global_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// A peer that does not show up is a dead or wedged process: after `timeout_ns` (minutes by default,
// TFC_COMM_TIMEOUT_S) the kernel traps, which surfaces as a launch failure on this rank.  The wait
// backs off (up to ~1 us between polls), so a long wait does not hammer NVLink or the peer's L2.
struct SpinWait {
  unsigned long long t0 = 0;
  unsigned ns = 32;
  __device__ __forceinline__ void pause(unsigned long long timeout_ns) {
    if (t0 == 0) t0 = global_ns();
    __nanosleep(ns);
    if (ns < 1024) ns <<= 1;
    else if (global_ns() - t0 > timeout_ns) __trap();
  }
};

__device__ __forceinline__ float* aux_of(uint32_t* pad) {
  return reinterpret_cast<float*>(pad + CB2_COMM_MAX_BLOCKS * CB2_MAX_RANKS);
}

// Block-level barrier with all peers' same-numbered block.  Threads [0, world) each handle one
// peer: publish my epoch into the peer's pad, then wait for the peer's epoch in mine.
__device__ __forceinline__ void peer_barrier(const DevCtx& c, uint32_t epoch) {
  __syncthreads();
  if (threadIdx.x < static_cast<unsigned>(c.world)) {
    const int peer = threadIdx.x;
    __threadfence_system();
    st_release_sys(c.pad[peer] + blockIdx.x * CB2_MAX_RANKS + c.rank, epoch);
    const uint32_t* mine = c.pad[c.rank] + blockIdx.x * CB2_MAX_RANKS + peer;
    SpinWait w;
    while (static_cast<int32_t>(ld_acquire_sys(mine) - epoch) < 0) w.pause(c.timeout_ns);
  }
  __syncthreads();
}

template <typename GT> __device__ __forceinline__ Vec8 mc_reduce8(const void* mc, size_t i);
template <> __device__ __forceinline__ Vec8 mc_reduce8<float>(const void* mc, size_t i) {
  const float* p = reinterpret_cast<const float*>(mc) + i;
  const uint4 a = mc_ld_reduce_f32x4(p), b = mc_ld_reduce_f32x4(p + 4);
  Vec8 r;
  r.v[0] = __uint_as_float(a.x); r.v[1] = __uint_as_float(a.y);
  r.v[2] = __uint_as_float(a.z); r.v[3] = __uint_as_float(a.w);
  r.v[4] = __uint_as_float(b.x); r.v[5] = __uint_as_float(b.y);
  r.v[6] = __uint_as_float(b.z); r.v[7] = __uint_as_float(b.w);
  return r;
}
template <> __device__ __forceinline__ Vec8 mc_reduce8<__nv_bfloat16>(const void* mc, size_t i) {
  return unpack8_bf16(mc_ld_reduce_bf16x8(reinterpret_cast<const __nv_bfloat16*>(mc) + i));
}
template <typename PT> __device__ __forceinline__ void mc_store8(void* mc, size_t i, const Vec8& r);
template <> __device__ __forceinline__ void mc_store8<float>(void* mc, size_t i, const Vec8& r) {
  float* p = reinterpret_cast<float*>(mc) + i;
  uint4 a, b;
  a.x = __float_as_uint(r.v[0]); a.y = __float_as_uint(r.v[1]);
  a.z = __float_as_uint(r.v[2]); a.w = __float_as_uint(r.v[3]);
  b.x = __float_as_uint(r.v[4]); b.y = __float_as_uint(r.v[5]);
  b.z = __float_as_uint(r.v[6]); b.w = __float_as_uint(r.v[7]);
  mc_st_v4(p, a);
  mc_st_v4(p + 4, b);
}
template <> __device__ __forceinline__ void mc_store8<__nv_bfloat16>(void* mc, size_t i, const Vec8& r) {
  mc_st_v4(reinterpret_cast<__nv_bfloat16*>(mc) + i, pack8_bf16(r));
}

struct FusedArgs {
  size_t off, len, state_off;
  float* master;
  float* s1;
  float* s2;
  const float* hyper;
  int zero_grads;
  int aux_len;
  float* aux_out;
};

// ---------------------------------------------------------------------------------------------
// Fused all-reduce + optimizer.  The unit of work is ONE 16-byte chunk (4 fp32 or 8 bf16): every
// load/store instruction of a warp then covers 512 contiguous bytes.  (A first version moved 8
// fp32 = two half-used 32-byte sectors per thread; peer memory bypasses the local L2, so each half
// sector crossed NVLink twice and fp32 bus bandwidth stalled at ~290 GB/s.)
// ---------------------------------------------------------------------------------------------
template <int N> struct VecN { float v[N]; };

template <typename T> __device__ __forceinline__ VecN<16 / sizeof(T)> unpack16(const uint4& w);
template <> __device__ __forceinline__ VecN<4> unpack16<float>(const uint4& w) {
  VecN<4> r;
  r.v[0] = __uint_as_float(w.x); r.v[1] = __uint_as_float(w.y);
  r.v[2] = __uint_as_float(w.z); r.v[3] = __uint_as_float(w.w);
  return r;
}
template <> __device__ __forceinline__ VecN<8> unpack16<__nv_bfloat16>(const uint4& w) {
  VecN<8> r;
  r.v[0] = bf16_lo(w.x); r.v[1] = bf16_hi(w.x); r.v[2] = bf16_lo(w.y); r.v[3] = bf16_hi(w.y);
  r.v[4] = bf16_lo(w.z); r.v[5] = bf16_hi(w.z); r.v[6] = bf16_lo(w.w); r.v[7] = bf16_hi(w.w);
  return r;
}
__device__ __forceinline__ uint4 pack16(const VecN<4>& r) {
  return make_uint4(__float_as_uint(r.v[0]), __float_as_uint(r.v[1]), __float_as_uint(r.v[2]),
                    __float_as_uint(r.v[3]));
}
__device__ __forceinline__ uint4 pack16(const VecN<8>& r) {
  return make_uint4(pack_bf16(r.v[0], r.v[1]), pack_bf16(r.v[2], r.v[3]), pack_bf16(r.v[4], r.v[5]),
                    pack_bf16(r.v[6], r.v[7]));
}
template <int N> __device__ __forceinline__ VecN<N> load_state(const float* p) {
  VecN<N> r;
#pragma unroll
  for (int k = 0; k < N; k += 4) {
    const float4 a = *reinterpret_cast<const float4*>(p + k);
    r.v[k] = a.x; r.v[k + 1] = a.y; r.v[k + 2] = a.z; r.v[k + 3] = a.w;
  }
  return r;
}
template <int N> __device__ __forceinline__ void store_state(float* p, const VecN<N>& r) {
#pragma unroll
  for (int k = 0; k < N; k += 4)
    *reinterpret_cast<float4*>(p + k) = make_float4(r.v[k], r.v[k + 1], r.v[k + 2], r.v[k + 3]);
}

template <int OPT, int N>
__device__ __forceinline__ VecN<N> updateN(const VecN<N>& p, const VecN<N>& g, float* s1, float* s2,
                                           size_t si, const Hyper& h) {
  VecN<N> a, b, out;
  if (OptTraits<OPT>::kS1) a = load_state<N>(s1 + si);
  if (OptTraits<OPT>::kS2) b = load_state<N>(s2 + si);
#pragma unroll
  for (int j = 0; j < N; ++j) out.v[j] = opt_update<OPT>(p.v[j], g.v[j], a.v[j], b.v[j], h);
  if (OptTraits<OPT>::kS1) store_state<N>(s1 + si, a);
  if (OptTraits<OPT>::kS2) store_state<N>(s2 + si, b);
  return out;
}

// Runtime optimizer dispatch (warp-uniform switch) keeps the number of kernel instantiations at
// schedule x dtype x world instead of multiplying by six optimizers.
template <int N>
__device__ __forceinline__ VecN<N> updateN_dyn(int opt, const VecN<N>& p, const VecN<N>& g,
                                               float* s1, float* s2, size_t si, const Hyper& h) {
  switch (opt) {
    case CB2_OPT_ADAM_KERAS: return updateN<CB2_OPT_ADAM_KERAS, N>(p, g, s1, s2, si, h);
    case CB2_OPT_ADAM_TORCH: return updateN<CB2_OPT_ADAM_TORCH, N>(p, g, s1, s2, si, h);
    case CB2_OPT_ADAMW: return updateN<CB2_OPT_ADAMW, N>(p, g, s1, s2, si, h);
    case CB2_OPT_SGD: return updateN<CB2_OPT_SGD, N>(p, g, s1, s2, si, h);
    case CB2_OPT_RMSPROP: return updateN<CB2_OPT_RMSPROP, N>(p, g, s1, s2, si, h);
    default: return updateN<CB2_OPT_NONE, N>(p, g, s1, s2, si, h);
  }
}

template <typename T> __device__ __forceinline__ uint4 mc_ld16(const void* mc);
template <> __device__ __forceinline__ uint4 mc_ld16<float>(const void* mc) { return mc_ld_reduce_f32x4(mc); }
template <> __device__ __forceinline__ uint4 mc_ld16<__nv_bfloat16>(const void* mc) { return mc_ld_reduce_bf16x8(mc); }

// WT = compile-time world size (2, 4, 8) or 0 for "any" (runtime loop, one chunk per iteration).
// Loads of a batch are all issued before the first use, so a thread keeps U x W 16-byte requests
// in flight across the ~2 us NVLink round trip.
template <int ALGO, typename T, int WT>
__global__ void __launch_bounds__(512)
comm_fused_kernel(const DevCtx c, const FusedArgs a, const int opt) {
  constexpr int N = 16 / sizeof(T);
  constexpr int WU = WT ? WT : 1;
  constexpr int U = ALGO == CB2_ALGO_NVLS ? 8 : (WT ? (16 / WT > 0 ? 16 / WT : 1) : 1);
  const uint32_t base = c.epochs[blockIdx.x];
  const Hyper h = load_hyper(a.hyper);
  const int W = WT ? WT : c.world;
  const int R = c.rank;
  const size_t tid = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x;
  const size_t nthreads = static_cast<size_t>(gridDim.x) * blockDim.x;

  // 1. every rank's gradients (and aux scalars) for this bucket are final
  if (W > 1) peer_barrier(c, base + 1);

  const bool oneshot = ALGO == CB2_ALGO_ONESHOT;
  const size_t span = oneshot ? a.len : a.len / W;     // elements this rank reduces
  const size_t nvec = span / N;
  const size_t ibase = oneshot ? a.off : a.off + static_cast<size_t>(R) * span;
  T* const my_param = reinterpret_cast<T*>(c.param[R]);

  for (size_t v0 = tid; v0 < nvec; v0 += nthreads * U) {
    uint4 raw[U][WU];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t v = v0 + static_cast<size_t>(u) * nthreads;
      if (v < nvec) {
        const size_t i = ibase + v * N;
        if (ALGO == CB2_ALGO_NVLS) {
          raw[u][0] = mc_ld16<T>(reinterpret_cast<const T*>(c.mc_grad) + i);
        } else if (WT) {
#pragma unroll
          for (int d = 0; d < WU; ++d) {
            // one-shot: fixed rank order => bitwise identical sums on every replica;
            // two-shot: start at my own arena and walk the ring so W readers hit W sources
            const int p = oneshot ? d : ((R + d) % WU);
            raw[u][d] = ld_stream_v4(reinterpret_cast<const T*>(c.grad[p]) + i);
          }
        } else {
          raw[u][0] = ld_stream_v4(reinterpret_cast<const T*>(c.grad[oneshot ? 0 : R]) + i);
        }
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t v = v0 + static_cast<size_t>(u) * nthreads;
      if (v >= nvec) continue;
      const size_t i = ibase + v * N;
      VecN<N> g = unpack16<T>(raw[u][0]);
      if (ALGO != CB2_ALGO_NVLS) {
        if (WT) {
#pragma unroll
          for (int d = 1; d < WU; ++d) {
            const VecN<N> o = unpack16<T>(raw[u][d]);
#pragma unroll
            for (int j = 0; j < N; ++j) g.v[j] += o.v[j];
          }
        } else {
          for (int d = 1; d < W; ++d) {
            int p = oneshot ? d : R + d;
            if (p >= W) p -= W;
            const VecN<N> o = unpack16<T>(ld_stream_v4(reinterpret_cast<const T*>(c.grad[p]) + i));
#pragma unroll
            for (int j = 0; j < N; ++j) g.v[j] += o.v[j];
          }
        }
      }
      const size_t si = a.state_off + v * N;
      VecN<N> pin;
      if (opt != CB2_OPT_NONE) {
        if (a.master) pin = load_state<N>(a.master + si);
        else pin = unpack16<T>(ld_stream_v4(my_param + i));
      }
      const VecN<N> pn = updateN_dyn<N>(opt, pin, g, a.s1, a.s2, si, h);
      if (opt != CB2_OPT_NONE && a.master) store_state<N>(a.master + si, pn);
      const uint4 packed = pack16(pn);
      if (oneshot) {
        st_stream_v4(my_param + i, packed);
      } else if (ALGO == CB2_ALGO_NVLS) {
        mc_st_v4(reinterpret_cast<T*>(c.mc_param) + i, packed);
      } else {
        for (int d = 0; d < W; ++d) {
          int p = R + d;
          if (p >= W) p -= W;
          st_stream_v4(reinterpret_cast<T*>(c.param[p]) + i, packed);
        }
      }
    }
  }

  // aux scalars (loss sum, hit count, ...) ride along: one-shot sum by block 0
  if (blockIdx.x == 0 && static_cast<int>(threadIdx.x) < a.aux_len) {
    float s = 0.f;
    for (int p = 0; p < W; ++p) {
      const float* src = aux_of(c.pad[p]) + threadIdx.x;
      s += __uint_as_float(ld_relaxed_sys(reinterpret_cast<const uint32_t*>(src)));
    }
    a.aux_out[threadIdx.x] = s;
  }

  // 2. all peers are done reading my gradients / aux and their parameter stores have landed
  if (W > 1) peer_barrier(c, base + 2);

  if (a.zero_grads) {
    // Block b zeroes exactly what the b-th blocks of all ranks have read, which is what the
    // barrier above covered: the b-th stripe of every shard (or of the whole range, one-shot).
    T* mine = reinterpret_cast<T*>(c.grad[R]);
    const int shards = oneshot ? 1 : W;
    const uint4 z = make_uint4(0, 0, 0, 0);
    for (int sidx = 0; sidx < shards; ++sidx)
      for (size_t v = tid; v < nvec; v += nthreads)
        st_stream_v4(mine + a.off + static_cast<size_t>(sidx) * span + v * N, z);
  }
  if (blockIdx.x == 0 && a.zero_grads && static_cast<int>(threadIdx.x) < a.aux_len)
    aux_of(c.pad[R])[threadIdx.x] = 0.f;

  if (threadIdx.x == 0) c.epochs[blockIdx.x] = base + 2;
}

// ---------------------------------------------------------------------------------------------
// LL ("low latency") schedules: flag-in-payload pushes, no barrier at all.
//
// A 16-byte gradient chunk travels as two 16-byte stores {d0, e, d1, e}, {d2, e, d3, e}: every
// 8-byte half carries the epoch e of the call, so the arrival of the data IS the signal - one
// NVLink traversal instead of "barrier, then load across the link".  The receiver polls its own
// memory.  Receive buffers alternate between two parities (e & 1): a rank can only be one call ahead
// of its slowest peer (it needs that peer's data of call e to finish call e), so data of call e + 2
// never lands on data of call e that has not been consumed.  Nobody reads a peer's gradient arena in
// these schedules, hence no closing barrier either: gradients are zeroed as soon as they are read.
//
//   LL1 (one-shot)  every rank pushes its whole bucket to every peer, sums all W copies in rank
//                   order (bitwise identical everywhere) and updates a replicated state.
//                   One hop; wire bytes 2 (W-1) S per rank: for buckets up to ~64 KiB.
//   LL2 (two-shot)  rank r owns shard r: peers push their part of it, r sums, updates its sharded
//                   state and pushes the new parameters back the same way.  Two hops; wire bytes
//                   4 (W-1)/W S per rank: for buckets up to a few MiB.
//
// Layout of a rank's LL buffer (uint4 units, n = chunk capacity of one slot):
//   [parity][region: 0 = gradients, 1 = parameters][sender rank][2 * n]   then the aux scalars:
//   [parity][sender][CB2_COMM_AUX_FLOATS] as 8-byte {value, e} pairs.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void ll_push(uint4* dst, const uint4& d, uint32_t e) {
  asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(dst), "r"(d.x), "r"(e), "r"(d.y), "r"(e)
               : "memory");
  asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(dst + 1), "r"(d.z), "r"(e), "r"(d.w),
               "r"(e)
               : "memory");
}
__device__ __forceinline__ uint4 ll_poll(const uint4* src, uint32_t e, unsigned long long timeout_ns) {
  uint4 a, b;
  SpinWait w;
  while (true) {
    asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(a.x), "=r"(a.y), "=r"(a.z), "=r"(a.w) : "l"(src) : "memory");
    asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(b.x), "=r"(b.y), "=r"(b.z), "=r"(b.w) : "l"(src + 1) : "memory");
    if (a.y == e && a.w == e && b.y == e && b.w == e) break;
    w.pause(timeout_ns);
  }
  return make_uint4(a.x, a.z, b.x, b.z);
}

// One chunk from each of the WT-1 peers' slots: every load is issued before the first flag is
// examined (independent requests overlap), and only the slots that were not complete are read again.
template <int WT>
__device__ __forceinline__ void ll_poll_all(uint4 (&out)[WT], const uint4* const (&src)[WT], int self,
                                            uint32_t e, unsigned long long timeout_ns) {
  uint4 a[WT], b[WT];
  uint32_t pending = ((1u << WT) - 1u) & ~(1u << self);
  SpinWait w;
  while (true) {
#pragma unroll
    for (int p = 0; p < WT; ++p) {
      if (pending & (1u << p)) {
        asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];"
                     : "=r"(a[p].x), "=r"(a[p].y), "=r"(a[p].z), "=r"(a[p].w) : "l"(src[p]) : "memory");
        asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];"
                     : "=r"(b[p].x), "=r"(b[p].y), "=r"(b[p].z), "=r"(b[p].w) : "l"(src[p] + 1) : "memory");
      }
    }
#pragma unroll
    for (int p = 0; p < WT; ++p) {
      if ((pending & (1u << p)) && a[p].y == e && a[p].w == e && b[p].y == e && b[p].w == e) {
        out[p] = make_uint4(a[p].x, a[p].z, b[p].x, b[p].z);
        pending &= ~(1u << p);
      }
    }
    if (!pending) break;
    w.pause(timeout_ns);
  }
}

__host__ __device__ inline size_t ll_slot_chunks(size_t ll_bytes, int world) {
  // bytes = 2 parities * (2 regions * W slots * 2n uint4 + W * AUX 8-byte pairs)
  const size_t aux = static_cast<size_t>(world) * CB2_COMM_AUX_FLOATS * 8;
  const size_t per_parity = ll_bytes / 2;
  if (per_parity <= aux) return 0;
  return (per_parity - aux) / 16 / (2 * static_cast<size_t>(world) * 2);
}

struct LLView {
  size_t n;            // chunk capacity of a slot
  size_t parity_u4;    // uint4 per parity
  __device__ __forceinline__ uint4* slot(uint4* base, uint32_t e, int region, int sender, int world) const {
    return base + (e & 1) * parity_u4 + (static_cast<size_t>(region) * world + sender) * 2 * n;
  }
  __device__ __forceinline__ uint2* aux(uint4* base, uint32_t e, int sender, int world) const {
    uint4* p = base + (e & 1) * parity_u4 + static_cast<size_t>(2) * world * 2 * n;
    return reinterpret_cast<uint2*>(p) + static_cast<size_t>(sender) * CB2_COMM_AUX_FLOATS;
  }
};

// aux scalars ride with the call: pushed to every peer as {value, e}, summed in rank order
__device__ __forceinline__ void ll_aux(const DevCtx& c, const FusedArgs& a, const LLView& lv, uint32_t e,
                                       int W, int R) {
  if (blockIdx.x != 0 || static_cast<int>(threadIdx.x) >= a.aux_len) return;
  float* mine = aux_of(c.pad[R]) + threadIdx.x;
  const float v = *mine;
  for (int p = 0; p < W; ++p) {
    if (p == R) continue;
    uint2* dst = lv.aux(c.ll[p], e, R, W) + threadIdx.x;
    asm volatile("st.volatile.global.v2.u32 [%0], {%1, %2};" ::"l"(dst), "r"(__float_as_uint(v)), "r"(e)
                 : "memory");
  }
  float s = 0.f;
  for (int p = 0; p < W; ++p) {
    if (p == R) { s += v; continue; }
    const uint2* src = lv.aux(c.ll[R], e, p, W) + threadIdx.x;
    uint32_t x, f;
    SpinWait w;
    while (true) {
      asm volatile("ld.volatile.global.v2.u32 {%0, %1}, [%2];" : "=r"(x), "=r"(f) : "l"(src) : "memory");
      if (f == e) break;
      w.pause(c.timeout_ns);
    }
    s += __uint_as_float(x);
  }
  a.aux_out[threadIdx.x] = s;
  if (a.zero_grads) *mine = 0.f;
}

// the call counter advances when the last block of the grid is done (no block of this launch can
// still read it then; the next launch on the stream sees the new value)
__device__ __forceinline__ void ll_finish(const DevCtx& c, uint32_t e) {
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    if (atomicAdd(c.ll_state + 1, 1u) == gridDim.x - 1) {
      c.ll_state[1] = 0;
      c.ll_state[0] = e;
    }
  }
}

template <typename T, int WT>
__global__ void __launch_bounds__(256)
comm_ll1_kernel(const DevCtx c, const FusedArgs a, const int opt) {
  constexpr int N = 16 / sizeof(T);
  const int W = WT, R = c.rank;
  const uint32_t e = *reinterpret_cast<volatile uint32_t*>(c.ll_state) + 1;
  const Hyper h = load_hyper(a.hyper);
  LLView lv;
  lv.n = ll_slot_chunks(c.ll_bytes, W);
  lv.parity_u4 = c.ll_bytes / 32;
  const size_t nvec = a.len / N;
  const size_t tid = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x;
  const size_t nthreads = static_cast<size_t>(gridDim.x) * blockDim.x;
  T* const my_grad = reinterpret_cast<T*>(c.grad[R]);
  T* const my_param = reinterpret_cast<T*>(c.param[R]);
  const uint4 zero = make_uint4(0, 0, 0, 0);

  for (size_t v = tid; v < nvec; v += nthreads) {
    uint4 part[WT];
    part[R] = ld_stream_v4(my_grad + a.off + v * N);
#pragma unroll
    for (int d = 1; d < WT; ++d) {
      const int p = (R + d) % WT;
      ll_push(lv.slot(c.ll[p], e, 0, R, W) + 2 * v, part[R], e);
    }
    if (a.zero_grads) st_stream_v4(my_grad + a.off + v * N, zero);
    const uint4* src[WT];
#pragma unroll
    for (int p = 0; p < WT; ++p) src[p] = lv.slot(c.ll[R], e, 0, p, W) + 2 * v;
    ll_poll_all<WT>(part, src, R, e, c.timeout_ns);
    VecN<N> g;
#pragma unroll
    for (int j = 0; j < N; ++j) g.v[j] = 0.f;
#pragma unroll
    for (int p = 0; p < WT; ++p) {            // rank order: the same sum on every replica
      const VecN<N> o = unpack16<T>(part[p]);
#pragma unroll
      for (int j = 0; j < N; ++j) g.v[j] += o.v[j];
    }
    const size_t i = a.off + v * N;
    const size_t si = a.state_off + v * N;
    VecN<N> pin;
    if (opt != CB2_OPT_NONE) {
      if (a.master) pin = load_state<N>(a.master + si);
      else pin = unpack16<T>(ld_stream_v4(my_param + i));
    }
    const VecN<N> pn = updateN_dyn<N>(opt, pin, g, a.s1, a.s2, si, h);
    if (opt != CB2_OPT_NONE && a.master) store_state<N>(a.master + si, pn);
    st_stream_v4(my_param + i, pack16(pn));
  }
  ll_aux(c, a, lv, e, W, R);
  ll_finish(c, e);
}

template <typename T, int WT>
__global__ void __launch_bounds__(256)
comm_ll2_kernel(const DevCtx c, const FusedArgs a, const int opt) {
  constexpr int N = 16 / sizeof(T);
  const int W = WT, R = c.rank;
  const uint32_t e = *reinterpret_cast<volatile uint32_t*>(c.ll_state) + 1;
  const Hyper h = load_hyper(a.hyper);
  LLView lv;
  lv.n = ll_slot_chunks(c.ll_bytes, W);
  lv.parity_u4 = c.ll_bytes / 32;
  const size_t span = a.len / W;             // elements per shard
  const size_t nvec = span / N;
  const size_t tid = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x;
  const size_t nthreads = static_cast<size_t>(gridDim.x) * blockDim.x;
  T* const my_grad = reinterpret_cast<T*>(c.grad[R]);
  T* const my_param = reinterpret_cast<T*>(c.param[R]);
  const uint4 zero = make_uint4(0, 0, 0, 0);

  for (size_t v = tid; v < nvec; v += nthreads) {
    // 1. my part of every peer's shard goes to its owner; my own shard stays in registers
    uint4 mine[WT];
#pragma unroll
    for (int p = 0; p < WT; ++p)
      mine[p] = ld_stream_v4(my_grad + a.off + static_cast<size_t>(p) * span + v * N);
#pragma unroll
    for (int d = 1; d < WT; ++d) {
      const int p = (R + d) % WT;
      ll_push(lv.slot(c.ll[p], e, 0, R, W) + 2 * v, mine[p], e);
    }
    if (a.zero_grads) {
#pragma unroll
      for (int p = 0; p < WT; ++p)
        st_stream_v4(my_grad + a.off + static_cast<size_t>(p) * span + v * N, zero);
    }
    // 2. sum my shard in rank order, update it, publish the new parameters
    uint4 part[WT];
    part[R] = mine[R];
    const uint4* src[WT];
#pragma unroll
    for (int p = 0; p < WT; ++p) src[p] = lv.slot(c.ll[R], e, 0, p, W) + 2 * v;
    ll_poll_all<WT>(part, src, R, e, c.timeout_ns);
    VecN<N> g;
#pragma unroll
    for (int j = 0; j < N; ++j) g.v[j] = 0.f;
#pragma unroll
    for (int p = 0; p < WT; ++p) {
      const VecN<N> o = unpack16<T>(part[p]);
#pragma unroll
      for (int j = 0; j < N; ++j) g.v[j] += o.v[j];
    }
    const size_t i = a.off + static_cast<size_t>(R) * span + v * N;
    const size_t si = a.state_off + v * N;
    VecN<N> pin;
    if (opt != CB2_OPT_NONE) {
      if (a.master) pin = load_state<N>(a.master + si);
      else pin = unpack16<T>(ld_stream_v4(my_param + i));
    }
    const VecN<N> pn = updateN_dyn<N>(opt, pin, g, a.s1, a.s2, si, h);
    if (opt != CB2_OPT_NONE && a.master) store_state<N>(a.master + si, pn);
    const uint4 packed = pack16(pn);
    st_stream_v4(my_param + i, packed);
#pragma unroll
    for (int d = 1; d < WT; ++d) {
      const int p = (R + d) % WT;
      ll_push(lv.slot(c.ll[p], e, 1, R, W) + 2 * v, packed, e);
    }
    // 3. the other shards' new parameters arrive the same way
#pragma unroll
    for (int p = 0; p < WT; ++p) src[p] = lv.slot(c.ll[R], e, 1, p, W) + 2 * v;
    ll_poll_all<WT>(part, src, R, e, c.timeout_ns);
#pragma unroll
    for (int p = 0; p < WT; ++p)
      if (p != R) st_stream_v4(my_param + a.off + static_cast<size_t>(p) * span + v * N, part[p]);
  }
  ll_aux(c, a, lv, e, W, R);
  ll_finish(c, e);
}

template <typename T>
__global__ void __launch_bounds__(512)
comm_broadcast_kernel(const DevCtx c, int root, size_t off, size_t len, int use_mc) {
  const uint32_t base = c.epochs[blockIdx.x];
  const int W = c.world;
  if (W > 1) peer_barrier(c, base + 1);  // receivers are ready to be overwritten
  if (c.rank == root) {
    const size_t nvec = len / 8;
    const size_t tid = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x;
    const size_t nthreads = static_cast<size_t>(gridDim.x) * blockDim.x;
    for (size_t vi = tid; vi < nvec; vi += nthreads) {
      const size_t i = off + vi * 8;
      const Vec8 v = load8<T>(reinterpret_cast<const T*>(c.param[root]), i);
      if (use_mc) {
        mc_store8<T>(c.mc_param, i, v);
      } else {
        for (int p = 0; p < W; ++p)
          if (p != root) store8<T>(reinterpret_cast<T*>(c.param[p]), i, v);
      }
    }
  }
  if (W > 1) peer_barrier(c, base + 2);  // root's stores are visible everywhere
  if (threadIdx.x == 0) c.epochs[blockIdx.x] = base + 2;
}

__global__ void comm_barrier_kernel(const DevCtx c) {
  const uint32_t base = c.epochs[blockIdx.x];
  if (c.world > 1) peer_barrier(c, base + 1);
  if (threadIdx.x == 0) c.epochs[blockIdx.x] = base + 1;
}

static int to_dev_ctx(const cb2_comm_ctx* ctx, DevCtx* d) {
  if (!ctx || ctx->world < 1 || ctx->world > CB2_MAX_RANKS || ctx->rank < 0 ||
      ctx->rank >= ctx->world || !ctx->epochs) {
    set_last_error("comm: invalid context");
    return CB2_ERR_ARG;
  }
  d->rank = ctx->rank;
  d->world = ctx->world;
  for (int i = 0; i < CB2_MAX_RANKS; ++i) {
    d->grad[i] = ctx->grad_ptrs[i];
    d->param[i] = ctx->param_ptrs[i];
    d->pad[i] = reinterpret_cast<uint32_t*>(ctx->pad_ptrs[i]);
  }
  d->mc_grad = ctx->mc_grad;
  d->mc_param = ctx->mc_param;
  d->epochs = ctx->epochs;
  for (int i = 0; i < CB2_MAX_RANKS; ++i) d->ll[i] = reinterpret_cast<uint4*>(ctx->ll_ptrs[i]);
  d->ll_bytes = ctx->ll_bytes;
  d->ll_state = ctx->ll_state;
  static const unsigned long long timeout_ns = [] {
    const char* e = getenv("TFC_COMM_TIMEOUT_S");
    const double sec = e ? atof(e) : 600.0;
    return static_cast<unsigned long long>((sec > 0 ? sec : 600.0) * 1e9);
  }();
  d->timeout_ns = timeout_ns;
  return 0;
}

template <int ALGO, typename T>
static int launch_ll(const DevCtx& d, const FusedArgs& a, int opt, int blocks, cudaStream_t st) {
#define CB2_LL(WT)                                                                    \
  if (ALGO == CB2_ALGO_LL1) comm_ll1_kernel<T, WT><<<blocks, 256, 0, st>>>(d, a, opt); \
  else comm_ll2_kernel<T, WT><<<blocks, 256, 0, st>>>(d, a, opt);
  switch (d.world) {
    case 2: CB2_LL(2) break;
    case 4: CB2_LL(4) break;
    case 8: CB2_LL(8) break;
    default:
      set_last_error("comm_fused: the LL schedules are built for 2, 4 or 8 ranks (world=%d)", d.world);
      return CB2_ERR_UNSUPPORTED;
  }
#undef CB2_LL
  return check_cuda(cudaGetLastError(), "comm_fused (LL) launch");
}

template <int ALGO, int WT>
static int launch_fused(const DevCtx& d, const FusedArgs& a, int opt, int gdt, int pdt, int blocks,
                        cudaStream_t st) {
  if (gdt != pdt) {
    set_last_error("comm_fused: gradient and parameter arenas must have the same dtype");
    return CB2_ERR_UNSUPPORTED;
  }
  if (gdt == CB2_DTYPE_F32) comm_fused_kernel<ALGO, float, WT><<<blocks, 512, 0, st>>>(d, a, opt);
  else comm_fused_kernel<ALGO, __nv_bfloat16, WT><<<blocks, 512, 0, st>>>(d, a, opt);
  return check_cuda(cudaGetLastError(), "comm_fused launch");
}

template <int ALGO>
static int dispatch_world(int opt, const DevCtx& d, const FusedArgs& a, int gdt, int pdt, int blocks,
                          cudaStream_t st) {
  if (opt < CB2_OPT_NONE || opt > CB2_OPT_RMSPROP) {
    set_last_error("comm_fused: unknown optimizer %d", opt);
    return CB2_ERR_ARG;
  }
  if (ALGO == CB2_ALGO_NVLS) return launch_fused<ALGO, 0>(d, a, opt, gdt, pdt, blocks, st);
  switch (d.world) {
    case 2: return launch_fused<ALGO, 2>(d, a, opt, gdt, pdt, blocks, st);
    case 4: return launch_fused<ALGO, 4>(d, a, opt, gdt, pdt, blocks, st);
    case 8: return launch_fused<ALGO, 8>(d, a, opt, gdt, pdt, blocks, st);
    default: return launch_fused<ALGO, 0>(d, a, opt, gdt, pdt, blocks, st);
  }
}

static int max_blocks_env() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("TFC_COMM_MAX_BLOCKS");
    v = e ? atoi(e) : 0;
    if (v < 0) v = 0;
    if (v > CB2_COMM_MAX_BLOCKS) v = CB2_COMM_MAX_BLOCKS;
  }
  return v;
}

}  // namespace cb2

using namespace cb2;

extern "C" size_t cb2_comm_ll_bytes(int world, size_t slot_payload) {
  const size_t chunks = (slot_payload + 15) / 16;
  const size_t per_parity = 2 * static_cast<size_t>(world) * 2 * chunks * 16 +
                            static_cast<size_t>(world) * CB2_COMM_AUX_FLOATS * 8;
  return 2 * per_parity;
}

extern "C" int cb2_comm_fused(const cb2_comm_ctx* ctx, int algo, int opt, int grad_dtype,
                              int param_dtype, size_t off, size_t len, float* master, float* s1,
                              float* s2, size_t state_off, const float* hyper, int zero_grads,
                              int aux_len, float* aux_out, int blocks, void* stream) {
  DevCtx d;
  int rc = to_dev_ctx(ctx, &d);
  if (rc) return rc;
  if (len == 0 && aux_len == 0) return 0;
  if ((off % 8) || (len % (8 * static_cast<size_t>(d.world)))) {
    set_last_error("comm_fused: off (%zu) must be a multiple of 8 and len (%zu) of 8*world", off, len);
    return CB2_ERR_ARG;
  }
  if (aux_len < 0 || aux_len > CB2_COMM_AUX_FLOATS || (aux_len > 0 && !aux_out)) {
    set_last_error("comm_fused: bad aux_len %d", aux_len);
    return CB2_ERR_ARG;
  }
  if (algo == CB2_ALGO_NVLS && (!d.mc_grad || !d.mc_param)) {
    set_last_error("comm_fused: NVLS requested but no multicast mapping in the context");
    return CB2_ERR_UNSUPPORTED;
  }
  if (opt != CB2_OPT_NONE && !master && param_dtype != CB2_DTYPE_F32) {
    set_last_error("comm_fused: bf16 parameters need an fp32 master copy");
    return CB2_ERR_ARG;
  }
  if (d.world == 1 && algo != CB2_ALGO_ONESHOT) algo = CB2_ALGO_TWOSHOT;
  if (algo == CB2_ALGO_LL1 || algo == CB2_ALGO_LL2) {
    if (grad_dtype != param_dtype) {
      set_last_error("comm_fused: gradient and parameter arenas must have the same dtype");
      return CB2_ERR_UNSUPPORTED;
    }
    const size_t cap = d.ll_state && d.ll[d.rank] ? ll_slot_chunks(d.ll_bytes, d.world) : 0;
    const size_t esz = grad_dtype == CB2_DTYPE_F32 ? 4 : 2;
    const size_t chunks = (algo == CB2_ALGO_LL1 ? len : len / d.world) * esz / 16;
    if (chunks > cap) {
      set_last_error("comm_fused: %zu bytes per slot exceed the LL buffer (%zu bytes per slot); use "
                     "TWOSHOT/NVLS for this bucket", chunks * 16, cap * 16);
      return CB2_ERR_UNSUPPORTED;
    }
    FusedArgs a;
    a.off = off; a.len = len; a.state_off = state_off;
    a.master = master; a.s1 = s1; a.s2 = s2; a.hyper = hyper;
    a.zero_grads = zero_grads; a.aux_len = aux_len; a.aux_out = aux_out;
    if (blocks <= 0) {
      // latency bound: one chunk per thread as long as the grid stays resident (every block waits
      // for peers, so all of them must be on the chip at once: 148 SMs, at least one 256-thread block each)
      size_t want = (chunks + 255) / 256;
      if (want > 128) want = 128;
      if (want < 1) want = 1;
      blocks = static_cast<int>(want);
    }
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    if (algo == CB2_ALGO_LL1)
      return grad_dtype == CB2_DTYPE_F32 ? launch_ll<CB2_ALGO_LL1, float>(d, a, opt, blocks, st)
                                         : launch_ll<CB2_ALGO_LL1, __nv_bfloat16>(d, a, opt, blocks, st);
    return grad_dtype == CB2_DTYPE_F32 ? launch_ll<CB2_ALGO_LL2, float>(d, a, opt, blocks, st)
                                       : launch_ll<CB2_ALGO_LL2, __nv_bfloat16>(d, a, opt, blocks, st);
  }
  if (blocks <= 0) {
    // enough 16-byte requests in flight to cover the ~2 us NVLink round trip at 900 GB/s, but
    // never more than the work: one 8-element vector per thread minimum
    const size_t work = (algo == CB2_ALGO_ONESHOT ? len : len / d.world) / 8;
    size_t want = (work + 511) / 512;
    size_t cap = algo == CB2_ALGO_NVLS ? 32 : 64;
    if (max_blocks_env()) cap = max_blocks_env();
    if (want > cap) want = cap;
    if (want < 1) want = 1;
    blocks = static_cast<int>(want);
  }
  if (blocks > CB2_COMM_MAX_BLOCKS) blocks = CB2_COMM_MAX_BLOCKS;
  FusedArgs a;
  a.off = off; a.len = len; a.state_off = state_off;
  a.master = master; a.s1 = s1; a.s2 = s2; a.hyper = hyper;
  a.zero_grads = zero_grads; a.aux_len = aux_len; a.aux_out = aux_out;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  switch (algo) {
    case CB2_ALGO_ONESHOT: return dispatch_world<CB2_ALGO_ONESHOT>(opt, d, a, grad_dtype, param_dtype, blocks, st);
    case CB2_ALGO_TWOSHOT: return dispatch_world<CB2_ALGO_TWOSHOT>(opt, d, a, grad_dtype, param_dtype, blocks, st);
    case CB2_ALGO_NVLS: return dispatch_world<CB2_ALGO_NVLS>(opt, d, a, grad_dtype, param_dtype, blocks, st);
  }
  set_last_error("comm_fused: unknown algorithm %d", algo);
  return CB2_ERR_ARG;
}

extern "C" int cb2_comm_broadcast(const cb2_comm_ctx* ctx, int root, int dtype, size_t off,
                                  size_t len, int use_multicast, int blocks, void* stream) {
  DevCtx d;
  int rc = to_dev_ctx(ctx, &d);
  if (rc) return rc;
  if (root < 0 || root >= d.world || (off % 8) || (len % 8)) {
    set_last_error("comm_broadcast: bad root/off/len");
    return CB2_ERR_ARG;
  }
  if (use_multicast && !d.mc_param) use_multicast = 0;
  if (blocks <= 0) {
    size_t want = (len / 8 + 511) / 512;
    if (want > 32) want = 32;
    if (want < 1) want = 1;
    blocks = static_cast<int>(want);
  }
  if (blocks > CB2_COMM_MAX_BLOCKS) blocks = CB2_COMM_MAX_BLOCKS;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (dtype == CB2_DTYPE_F32)
    comm_broadcast_kernel<float><<<blocks, 512, 0, st>>>(d, root, off, len, use_multicast);
  else
    comm_broadcast_kernel<__nv_bfloat16><<<blocks, 512, 0, st>>>(d, root, off, len, use_multicast);
  return check_cuda(cudaGetLastError(), "comm_broadcast launch");
}

extern "C" int cb2_comm_barrier(const cb2_comm_ctx* ctx, void* stream) {
  DevCtx d;
  int rc = to_dev_ctx(ctx, &d);
  if (rc) return rc;
  comm_barrier_kernel<<<1, 32, 0, reinterpret_cast<cudaStream_t>(stream)>>>(d);
  return check_cuda(cudaGetLastError(), "comm_barrier launch");
}
