// Host-memory twin of the peer-memory collectives (csrc/kernels/comm.cu), over POSIX shared memory.
//
// The reference has no collective code of its own: MirroredStrategy / MultiWorkerMirroredStrategy
// (src/python/tensorflow_cloud/core/preprocess.py:137-143) hand that to TensorFlow, and on CPU
// workers TensorFlow runs RING collectives (SURVEY.md sections 2.4 / 5.8).  This file is what CPU
// ranks of one box use here, and it is deliberately the SAME protocol as the GPU kernels so that the
// schedules and the flag protocol can be exercised - and put under ThreadSanitizer - without a GPU:
//
//   * every rank owns one arena of the segment ("symmetric memory"): rank r's arena is mapped by
//     every other rank, gradients are produced straight into it;
//   * a signal pad per rank holds one monotonic epoch counter per source rank; arrival is a
//     release store into every peer's pad, waiting is an acquire load of the own pad, every spin is
//     bounded by a watchdog (a dead peer makes the call fail with its rank, not hang);
//   * one-shot all-reduce: barrier, every rank sums all arenas in rank order (same order everywhere:
//     bit-identical results on all ranks), barrier, write back into its own arena;
//   * two-shot all-reduce: barrier, rank r sums shard r of all arenas, stores the result into shard
//     r of EVERY arena (the all-gather is done by peer stores), barrier;
//   * broadcast: root stores into every arena, barrier;
//   * a small vector of fp32 scalars (loss / accuracy / count of a step) rides in the pad and is
//     summed by the same barriers.
//
// fp32 and bf16 payloads; bf16 is accumulated in fp32 and rounded to nearest even once.
#include "rt_common.h"
#include "cloud_b200_runtime.h"   // the public declarations must match the definitions below

#include <atomic>
#include <errno.h>
#include <fcntl.h>
#include <new>
#include <sched.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <time.h>
#include <unistd.h>

namespace {

constexpr uint32_t kMagic = 0xCB2005A1u;
constexpr int kMaxRanks = 64;
constexpr int kAuxFloats = 256;
constexpr size_t kAlign = 4096;

struct alignas(64) Flag {
  std::atomic<uint32_t> epoch;
};

// One per rank, written by the peers (flags) and by the owner (aux).
struct Pad {
  Flag arrived[kMaxRanks];       // arrived[src]: last epoch rank `src` has reached
  float aux[kAuxFloats];
};

struct Header {
  uint32_t magic;
  int32_t world;
  uint64_t arena_bytes;
  std::atomic<int32_t> attached;
};

size_t round_up(size_t x, size_t a) { return (x + a - 1) / a * a; }
size_t pads_offset() { return round_up(sizeof(Header), kAlign); }
size_t arenas_offset(int world) { return pads_offset() + round_up(sizeof(Pad) * world, kAlign); }
size_t segment_bytes(int world, size_t arena_bytes) {
  return arenas_offset(world) + round_up(arena_bytes, kAlign) * world;
}

struct Comm {
  char* base = nullptr;
  size_t bytes = 0;
  bool owns_mapping = false;
  int rank = 0, world = 1;
  size_t arena_bytes = 0;
  uint32_t epoch = 0;            // this rank's count of barriers so far
  double timeout_s = 120.0;
  float* scratch = nullptr;      // fp32 accumulator for one call
  size_t scratch_elems = 0;

  Header* header() const { return reinterpret_cast<Header*>(base); }
  Pad* pad(int r) const { return reinterpret_cast<Pad*>(base + pads_offset()) + r; }
  char* arena(int r) const { return base + arenas_offset(world) + round_up(arena_bytes, kAlign) * r; }
};

double now_s() {
  timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return ts.tv_sec + 1e-9 * ts.tv_nsec;
}

// Everybody has reached barrier number `c->epoch + 1`.  The release store publishes every write
// this rank made before the barrier; the acquire loads make the peers' earlier writes visible.
int barrier(Comm* c) {
  const uint32_t e = ++c->epoch;
  for (int p = 0; p < c->world; ++p) c->pad(p)->arrived[c->rank].epoch.store(e, std::memory_order_release);
  Pad* mine = c->pad(c->rank);
  const double deadline = now_s() + c->timeout_s;
  for (int src = 0; src < c->world; ++src) {
    unsigned spins = 0;
    // epochs are monotonic; a peer may already be one barrier ahead, never behind after this loop
    while (static_cast<int32_t>(mine->arrived[src].epoch.load(std::memory_order_acquire) - e) < 0) {
      if ((++spins & 0x3ff) == 0) {
        if (now_s() > deadline) {
          cb2rt::set_error("shm collective: rank %d waited %.0f s at barrier %u for rank %d (at %u)",
                           c->rank, c->timeout_s, e, src,
                           mine->arrived[src].epoch.load(std::memory_order_relaxed));
          return -2;
        }
        sched_yield();
      }
    }
  }
  return 0;
}

inline float bf16_to_f32(uint16_t h) {
  uint32_t u = static_cast<uint32_t>(h) << 16;
  float f;
  memcpy(&f, &u, 4);
  return f;
}

inline uint16_t f32_to_bf16(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return static_cast<uint16_t>((u >> 16) | 0x40);   // NaN
  u += 0x7fffu + ((u >> 16) & 1u);                                                        // RNE
  return static_cast<uint16_t>(u >> 16);
}

int ensure_scratch(Comm* c, size_t elems) {
  if (elems == 0 || c->scratch_elems >= elems) return 0;
  free(c->scratch);
  c->scratch = static_cast<float*>(malloc(elems * sizeof(float)));
  c->scratch_elems = c->scratch ? elems : 0;
  if (!c->scratch) {
    cb2rt::set_error("shm collective: out of memory for %zu scratch floats", elems);
    return -1;
  }
  return 0;
}

// acc[0..n) = sum over ranks (in rank order) of arena_r[off + lo .. off + lo + n), times scale.
template <bool kBf16>
void reduce_range(Comm* c, size_t off_elems, size_t lo, size_t n, float scale, float* acc) {
  for (int r = 0; r < c->world; ++r) {
    const char* a = c->arena(r);
    if (kBf16) {
      const uint16_t* src = reinterpret_cast<const uint16_t*>(a) + off_elems + lo;
      if (r == 0) for (size_t i = 0; i < n; ++i) acc[i] = bf16_to_f32(src[i]);
      else for (size_t i = 0; i < n; ++i) acc[i] += bf16_to_f32(src[i]);
    } else {
      const float* src = reinterpret_cast<const float*>(a) + off_elems + lo;
      if (r == 0) memcpy(acc, src, n * sizeof(float));
      else for (size_t i = 0; i < n; ++i) acc[i] += src[i];
    }
  }
  if (scale != 1.0f) for (size_t i = 0; i < n; ++i) acc[i] *= scale;
}

template <bool kBf16>
void store_range(Comm* c, int dst_rank, size_t off_elems, size_t lo, size_t n, const float* acc) {
  char* a = c->arena(dst_rank);
  if (kBf16) {
    uint16_t* dst = reinterpret_cast<uint16_t*>(a) + off_elems + lo;
    for (size_t i = 0; i < n; ++i) dst[i] = f32_to_bf16(acc[i]);
  } else {
    memcpy(reinterpret_cast<float*>(a) + off_elems + lo, acc, n * sizeof(float));
  }
}

int reduce_aux(Comm* c, int aux_len, float* aux_out) {
  // called between two barriers: every pad's aux is stable
  for (int i = 0; i < aux_len; ++i) {
    float s = 0.f;
    for (int r = 0; r < c->world; ++r) s += c->pad(r)->aux[i];
    aux_out[i] = s;
  }
  return 0;
}

template <bool kBf16>
int allreduce(Comm* c, int algo, size_t off, size_t len, float scale, int aux_len, float* aux_out) {
  int rc = barrier(c);                                  // all gradients (and aux scalars) are written
  if (rc) return rc;
  if (algo == 0) {                                      // one-shot
    if ((rc = ensure_scratch(c, len))) return rc;
    reduce_range<kBf16>(c, off, 0, len, scale, c->scratch);
    if (aux_len) reduce_aux(c, aux_len, aux_out);
    if ((rc = barrier(c))) return rc;                   // nobody reads the inputs any more
    // from here on every rank touches its own arena and pad only; the first barrier of the next
    // call orders the next round of writes against these
    store_range<kBf16>(c, c->rank, off, 0, len, c->scratch);
    if (aux_len) memset(c->pad(c->rank)->aux, 0, sizeof(float) * aux_len);
    return 0;
  }
  // two-shot: reduce-scatter by peer loads, all-gather by peer stores
  const size_t shard = (len + c->world - 1) / c->world;
  const size_t lo = shard * c->rank < len ? shard * c->rank : len;
  const size_t n = lo + shard <= len ? shard : len - lo;
  if ((rc = ensure_scratch(c, n ? n : 1))) return rc;
  if (n) reduce_range<kBf16>(c, off, lo, n, scale, c->scratch);
  if (aux_len) reduce_aux(c, aux_len, aux_out);
  // shard r of every arena is read by rank r only, and rank r is done with it: no barrier needed
  // before overwriting it everywhere
  for (int p = 0; p < c->world && n; ++p) store_range<kBf16>(c, (c->rank + p) % c->world, off, lo, n, c->scratch);
  if ((rc = barrier(c))) return rc;                     // every shard has landed everywhere
  // peers read this pad's scalars before they arrived at the barrier above
  if (aux_len) memset(c->pad(c->rank)->aux, 0, sizeof(float) * aux_len);
  return 0;
}

}  // namespace

extern "C" {

size_t cb2rt_shm_segment_bytes(int world, size_t arena_bytes) { return segment_bytes(world, arena_bytes); }
int cb2rt_shm_max_ranks(void) { return kMaxRanks; }
int cb2rt_shm_aux_floats(void) { return kAuxFloats; }

// Formats a zero-filled region of `segment_bytes` as a segment (done once, by whoever created it).
int cb2rt_shm_format(void* base, int world, size_t arena_bytes) {
  if (world < 1 || world > kMaxRanks) {
    cb2rt::set_error("shm collective: world size %d outside 1..%d", world, kMaxRanks);
    return -1;
  }
  Header* h = new (base) Header;
  h->world = world;
  h->arena_bytes = arena_bytes;
  h->attached.store(0, std::memory_order_relaxed);
  Pad* pads = reinterpret_cast<Pad*>(static_cast<char*>(base) + pads_offset());
  for (int r = 0; r < world; ++r) new (pads + r) Pad();
  std::atomic_thread_fence(std::memory_order_release);
  h->magic = kMagic;
  return 0;
}

// Attaches rank `rank` to an already mapped and formatted segment (threads of one process sharing
// one mapping - the sanitizer tests - or the tail of cb2rt_shm_open).
void* cb2rt_shm_attach(void* base, int rank, double timeout_s) {
  Header* h = static_cast<Header*>(base);
  if (h->magic != kMagic) {
    cb2rt::set_error("shm collective: segment is not formatted");
    return nullptr;
  }
  if (rank < 0 || rank >= h->world) {
    cb2rt::set_error("shm collective: rank %d outside world %d", rank, h->world);
    return nullptr;
  }
  Comm* c = new (std::nothrow) Comm;
  if (!c) return nullptr;
  c->base = static_cast<char*>(base);
  c->rank = rank;
  c->world = h->world;
  c->arena_bytes = h->arena_bytes;
  c->bytes = segment_bytes(c->world, c->arena_bytes);
  if (timeout_s > 0) c->timeout_s = timeout_s;
  h->attached.fetch_add(1, std::memory_order_acq_rel);
  return c;
}

// Creates (`create` != 0: exclusive, stale segment of the same name replaced) or opens the named
// segment and attaches `rank`.  Openers must be started after the creator returned.
void* cb2rt_shm_open(const char* name, int rank, int world, size_t arena_bytes, int create, double timeout_s) {
  const size_t bytes = segment_bytes(world, arena_bytes);
  int fd;
  if (create) {
    shm_unlink(name);
    fd = shm_open(name, O_CREAT | O_EXCL | O_RDWR, 0600);
    if (fd >= 0 && ftruncate(fd, static_cast<off_t>(bytes)) != 0) {
      cb2rt::set_error("shm collective: ftruncate(%s, %zu): %s", name, bytes, strerror(errno));
      close(fd);
      shm_unlink(name);
      return nullptr;
    }
    // Reserve the pages now: a tmpfs that is too small (containers often mount 64 MB at /dev/shm)
    // must fail here with an error, not later with SIGBUS on the first touch of a missing page.
    if (fd >= 0) {
      const int rc = posix_fallocate(fd, 0, static_cast<off_t>(bytes));
      if (rc != 0) {
        cb2rt::set_error("shm collective: cannot reserve %zu bytes of shared memory for %s: %s", bytes, name,
                         strerror(rc));
        close(fd);
        shm_unlink(name);
        return nullptr;
      }
    }
  } else {
    fd = shm_open(name, O_RDWR, 0600);
  }
  if (fd < 0) {
    cb2rt::set_error("shm collective: shm_open(%s): %s", name, strerror(errno));
    return nullptr;
  }
  if (!create) {
    struct stat st;
    if (fstat(fd, &st) != 0 || static_cast<size_t>(st.st_size) < bytes) {
      cb2rt::set_error("shm collective: segment %s is smaller than world %d x %zu bytes needs", name, world,
                       arena_bytes);
      close(fd);
      return nullptr;
    }
  }
  void* base = mmap(nullptr, bytes, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
  close(fd);
  if (base == MAP_FAILED) {
    cb2rt::set_error("shm collective: mmap(%s, %zu): %s", name, bytes, strerror(errno));
    return nullptr;
  }
  if (create && cb2rt_shm_format(base, world, arena_bytes) != 0) {
    munmap(base, bytes);
    return nullptr;
  }
  Comm* c = static_cast<Comm*>(cb2rt_shm_attach(base, rank, timeout_s));
  if (!c) {
    munmap(base, bytes);
    return nullptr;
  }
  if (c->world != world || c->arena_bytes != arena_bytes) {
    cb2rt::set_error("shm collective: segment %s was created for world %d x %zu bytes", name, c->world,
                     c->arena_bytes);
    delete c;
    munmap(base, bytes);
    return nullptr;
  }
  c->owns_mapping = true;
  return c;
}

int cb2rt_shm_unlink(const char* name) { return shm_unlink(name); }

void cb2rt_shm_close(void* handle) {
  Comm* c = static_cast<Comm*>(handle);
  if (!c) return;
  free(c->scratch);
  if (c->owns_mapping) munmap(c->base, c->bytes);
  delete c;
}

int cb2rt_shm_rank(void* handle) { return static_cast<Comm*>(handle)->rank; }
int cb2rt_shm_world(void* handle) { return static_cast<Comm*>(handle)->world; }
int cb2rt_shm_attached(void* handle) {
  return static_cast<Comm*>(handle)->header()->attached.load(std::memory_order_acquire);
}
void* cb2rt_shm_arena(void* handle, int rank) {
  Comm* c = static_cast<Comm*>(handle);
  return (rank < 0 || rank >= c->world) ? nullptr : c->arena(rank);
}
float* cb2rt_shm_aux(void* handle) {
  Comm* c = static_cast<Comm*>(handle);
  return c->pad(c->rank)->aux;
}
void cb2rt_shm_set_timeout(void* handle, double seconds) { static_cast<Comm*>(handle)->timeout_s = seconds; }

int cb2rt_shm_barrier(void* handle) { return barrier(static_cast<Comm*>(handle)); }

// In-place sum over ranks of elements [off, off + len) of every arena, times `scale`.
// algo: 0 one-shot, 1 two-shot.  dtype: 0 fp32, 1 bf16.  With aux_len > 0 the first aux_len scalars
// of every rank's pad are summed into aux_out and the pad is cleared.  Collective: every rank calls
// it with the same arguments.  Returns 0, -1 (argument / memory) or -2 (a peer did not arrive).
int cb2rt_shm_allreduce(void* handle, int algo, int dtype, size_t off, size_t len, float scale, int aux_len,
                        float* aux_out) {
  Comm* c = static_cast<Comm*>(handle);
  const size_t item = dtype == 1 ? 2 : 4;
  if (algo < 0 || algo > 1 || dtype < 0 || dtype > 1 || aux_len < 0 || aux_len > kAuxFloats ||
      (aux_len && !aux_out)) {
    cb2rt::set_error("shm collective: bad arguments (algo %d, dtype %d, aux_len %d)", algo, dtype, aux_len);
    return -1;
  }
  if ((off + len) * item > c->arena_bytes) {
    cb2rt::set_error("shm collective: elements [%zu, %zu) of %zu bytes exceed the %zu-byte arena", off, off + len,
                     item, c->arena_bytes);
    return -1;
  }
  return dtype == 1 ? allreduce<true>(c, algo, off, len, scale, aux_len, aux_out)
                    : allreduce<false>(c, algo, off, len, scale, aux_len, aux_out);
}

// Root's bytes [off, off + nbytes) of its arena into the same place of every arena.
int cb2rt_shm_broadcast(void* handle, int root, size_t off, size_t nbytes) {
  Comm* c = static_cast<Comm*>(handle);
  if (root < 0 || root >= c->world || off + nbytes > c->arena_bytes) {
    cb2rt::set_error("shm collective: bad broadcast (root %d, bytes [%zu, %zu))", root, off, off + nbytes);
    return -1;
  }
  int rc = barrier(c);                                  // nobody still uses the destination
  if (rc) return rc;
  if (c->rank == root)
    for (int p = 0; p < c->world; ++p)
      if (p != root) memcpy(c->arena(p) + off, c->arena(root) + off, nbytes);
  return barrier(c);
}

}  // extern "C"
