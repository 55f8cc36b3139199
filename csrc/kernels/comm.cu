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
    const long long t0 = clock64();
    while (static_cast<int32_t>(ld_acquire_sys(mine) - epoch) < 0) {
      if (clock64() - t0 > kSpinTimeoutCycles) __trap();
    }
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

// Runtime optimizer dispatch (warp-uniform switch) keeps the number of kernel instantiations at
// schedule x dtypes x world instead of multiplying by six optimizers.
__device__ __forceinline__ Vec8 update8_dyn(int opt, const Vec8& p, const Vec8& g, float* s1,
                                            float* s2, size_t si, const Hyper& h) {
  switch (opt) {
    case CB2_OPT_ADAM_KERAS: return update8<CB2_OPT_ADAM_KERAS>(p, g, s1, s2, si, h);
    case CB2_OPT_ADAM_TORCH: return update8<CB2_OPT_ADAM_TORCH>(p, g, s1, s2, si, h);
    case CB2_OPT_ADAMW: return update8<CB2_OPT_ADAMW>(p, g, s1, s2, si, h);
    case CB2_OPT_SGD: return update8<CB2_OPT_SGD>(p, g, s1, s2, si, h);
    case CB2_OPT_RMSPROP: return update8<CB2_OPT_RMSPROP>(p, g, s1, s2, si, h);
    default: return update8<CB2_OPT_NONE>(p, g, s1, s2, si, h);
  }
}

// 8 elements as raw 16-byte chunks: loads are issued back to back and only unpacked afterwards,
// so a thread keeps U x W requests in flight across the ~2 us NVLink round trip.
template <typename GT> struct Raw8 { uint4 q[sizeof(GT) == 4 ? 2 : 1]; };

template <typename GT>
__device__ __forceinline__ Raw8<GT> ld_raw8(const void* base, size_t i) {
  Raw8<GT> r;
  const GT* p = reinterpret_cast<const GT*>(base) + i;
  r.q[0] = ld_stream_v4(p);
  if (sizeof(GT) == 4) r.q[sizeof(GT) == 4 ? 1 : 0] = ld_stream_v4(reinterpret_cast<const float*>(p) + 4);
  return r;
}
template <typename GT>
__device__ __forceinline__ Raw8<GT> mc_ld_raw8(const void* mc, size_t i) {
  Raw8<GT> r;
  const GT* p = reinterpret_cast<const GT*>(mc) + i;
  if (sizeof(GT) == 4) {
    r.q[0] = mc_ld_reduce_f32x4(p);
    r.q[sizeof(GT) == 4 ? 1 : 0] = mc_ld_reduce_f32x4(reinterpret_cast<const float*>(p) + 4);
  } else {
    r.q[0] = mc_ld_reduce_bf16x8(p);
  }
  return r;
}
template <typename GT>
__device__ __forceinline__ Vec8 unpack_raw8(const Raw8<GT>& r) {
  if (sizeof(GT) == 4) {
    Vec8 v;
    const uint4 a = r.q[0], b = r.q[sizeof(GT) == 4 ? 1 : 0];
    v.v[0] = __uint_as_float(a.x); v.v[1] = __uint_as_float(a.y);
    v.v[2] = __uint_as_float(a.z); v.v[3] = __uint_as_float(a.w);
    v.v[4] = __uint_as_float(b.x); v.v[5] = __uint_as_float(b.y);
    v.v[6] = __uint_as_float(b.z); v.v[7] = __uint_as_float(b.w);
    return v;
  }
  return unpack8_bf16(r.q[0]);
}

// WT = compile-time world size (2, 4, 8) or 0 for "any" (runtime loop, one vector per iteration).
template <int ALGO, typename GT, typename PT, int WT>
__global__ void __launch_bounds__(512)
comm_fused_kernel(const DevCtx c, const FusedArgs a, const int opt) {
  const uint32_t base = c.epochs[blockIdx.x];
  const Hyper h = load_hyper(a.hyper);
  const int W = WT ? WT : c.world;
  const int R = c.rank;
  const size_t tid = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x;
  const size_t nthreads = static_cast<size_t>(gridDim.x) * blockDim.x;
  constexpr int Q = sizeof(GT) == 4 ? 2 : 1;
  constexpr int WU = WT ? WT : 1;
  constexpr int U = ALGO == CB2_ALGO_NVLS ? 8 / Q : (WT ? (16 / (WT * Q) > 0 ? 16 / (WT * Q) : 1) : 1);

  // 1. every rank's gradients (and aux scalars) for this bucket are final
  if (W > 1) peer_barrier(c, base + 1);

  const bool oneshot = ALGO == CB2_ALGO_ONESHOT;
  const size_t span = oneshot ? a.len : a.len / W;     // elements this rank reduces
  const size_t nvec = span / 8;
  const size_t ibase = oneshot ? a.off : a.off + static_cast<size_t>(R) * span;

  for (size_t v0 = tid; v0 < nvec; v0 += nthreads * U) {
    Raw8<GT> raw[U][WU];
    // ---- issue every load of this batch ----------------------------------------------------
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t v = v0 + static_cast<size_t>(u) * nthreads;
      if (v < nvec) {
        const size_t i = ibase + v * 8;
        if (ALGO == CB2_ALGO_NVLS) {
          raw[u][0] = mc_ld_raw8<GT>(c.mc_grad, i);
        } else if (WT) {
#pragma unroll
          for (int d = 0; d < WU; ++d) {
            // one-shot: fixed rank order => bitwise identical sums on every replica;
            // two-shot: start at my own arena and walk the ring so W readers hit W sources
            const int p = oneshot ? d : ((R + d) % WU);
            raw[u][d] = ld_raw8<GT>(c.grad[p], i);
          }
        } else {
          raw[u][0] = ld_raw8<GT>(c.grad[oneshot ? 0 : R], i);
        }
      }
    }
    // ---- reduce, update, publish --------------------------------------------------------------
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t v = v0 + static_cast<size_t>(u) * nthreads;
      if (v >= nvec) continue;
      const size_t i = ibase + v * 8;
      Vec8 g = unpack_raw8<GT>(raw[u][0]);
      if (ALGO != CB2_ALGO_NVLS) {
        if (WT) {
#pragma unroll
          for (int d = 1; d < WU; ++d) {
            const Vec8 o = unpack_raw8<GT>(raw[u][d]);
#pragma unroll
            for (int j = 0; j < 8; ++j) g.v[j] += o.v[j];
          }
        } else {
          for (int d = 1; d < W; ++d) {
            int p = oneshot ? d : R + d;
            if (p >= W) p -= W;
            const Vec8 o = load8<GT>(reinterpret_cast<const GT*>(c.grad[p]), i);
#pragma unroll
            for (int j = 0; j < 8; ++j) g.v[j] += o.v[j];
          }
        }
      }
      const size_t si = a.state_off + v * 8;
      Vec8 pin;
      if (opt != CB2_OPT_NONE) {
        if (a.master) pin = load8_f32(a.master + si);
        else pin = load8<PT>(reinterpret_cast<const PT*>(c.param[R]), i);
      }
      const Vec8 pn = update8_dyn(opt, pin, g, a.s1, a.s2, si, h);
      if (opt != CB2_OPT_NONE && a.master) store8_f32(a.master + si, pn);
      if (oneshot) {
        store8<PT>(reinterpret_cast<PT*>(c.param[R]), i, pn);
      } else if (ALGO == CB2_ALGO_NVLS) {
        mc_store8<PT>(c.mc_param, i, pn);
      } else {
        for (int d = 0; d < W; ++d) {
          int p = R + d;
          if (p >= W) p -= W;
          store8<PT>(reinterpret_cast<PT*>(c.param[p]), i, pn);
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
    GT* mine = reinterpret_cast<GT*>(c.grad[R]);
    const int shards = oneshot ? 1 : W;
    for (int sidx = 0; sidx < shards; ++sidx)
      for (size_t v = tid; v < nvec; v += nthreads)
        zero8<GT>(mine, a.off + static_cast<size_t>(sidx) * span + v * 8);
  }
  if (blockIdx.x == 0 && a.zero_grads && static_cast<int>(threadIdx.x) < a.aux_len)
    aux_of(c.pad[R])[threadIdx.x] = 0.f;

  if (threadIdx.x == 0) c.epochs[blockIdx.x] = base + 2;
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
  return 0;
}

template <int ALGO, int WT>
static int launch_fused(const DevCtx& d, const FusedArgs& a, int opt, int gdt, int pdt, int blocks,
                        cudaStream_t st) {
#define CB2_LAUNCH(GT, PT) comm_fused_kernel<ALGO, GT, PT, WT><<<blocks, 512, 0, st>>>(d, a, opt)
  if (gdt == CB2_DTYPE_F32 && pdt == CB2_DTYPE_F32) CB2_LAUNCH(float, float);
  else if (gdt == CB2_DTYPE_BF16 && pdt == CB2_DTYPE_BF16) CB2_LAUNCH(__nv_bfloat16, __nv_bfloat16);
  else if (gdt == CB2_DTYPE_F32 && pdt == CB2_DTYPE_BF16) CB2_LAUNCH(float, __nv_bfloat16);
  else CB2_LAUNCH(__nv_bfloat16, float);
#undef CB2_LAUNCH
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
  if (d.world == 1 && algo == CB2_ALGO_NVLS) algo = CB2_ALGO_TWOSHOT;
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
