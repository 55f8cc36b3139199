// Unit test of the shared-memory collectives (shm_comm.cc): N threads attach as N ranks to ONE
// mapping of a segment, so the ThreadSanitizer build of this test checks the flag protocol itself -
// every read of a peer's arena must be ordered after that peer's writes by a release/acquire pair.
// Ranks arrive in random order with random delays (the skewed-arrival stress of SURVEY.md section
// 5.2), results are compared with a serial sum, and a missing rank must make the others fail with a
// watchdog error instead of hanging.  New tier: the reference has no collective code to test.
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <thread>
#include <vector>
#include <chrono>

extern "C" {
const char* cb2rt_last_error();
size_t cb2rt_shm_segment_bytes(int world, size_t arena_bytes);
int cb2rt_shm_format(void* base, int world, size_t arena_bytes);
void* cb2rt_shm_attach(void* base, int rank, double timeout_s);
void cb2rt_shm_close(void* handle);
void* cb2rt_shm_arena(void* handle, int rank);
float* cb2rt_shm_aux(void* handle);
int cb2rt_shm_barrier(void* handle);
int cb2rt_shm_allreduce(void* handle, int algo, int dtype, size_t off, size_t len, float scale, int aux_len,
                        float* aux_out);
int cb2rt_shm_broadcast(void* handle, int root, size_t off, size_t nbytes);
void* cb2rt_shm_open(const char* name, int rank, int world, size_t arena_bytes, int create, double timeout_s);
int cb2rt_shm_unlink(const char* name);
}

static int g_failed = 0;
#define CHECK(cond)                                                                            \
  do {                                                                                         \
    if (!(cond)) {                                                                             \
      std::fprintf(stderr, "CHECK failed %s:%d: %s (last error: %s)\n", __FILE__, __LINE__, #cond, \
                   cb2rt_last_error());                                                        \
      ++g_failed;                                                                              \
    }                                                                                          \
  } while (0)

static float bf16_round(float f) {
  uint32_t u;
  std::memcpy(&u, &f, 4);
  u += 0x7fffu + ((u >> 16) & 1u);
  u &= 0xffff0000u;
  std::memcpy(&f, &u, 4);
  return f;
}

static uint16_t to_bf16(float f) {
  uint32_t u;
  std::memcpy(&u, &f, 4);
  u += 0x7fffu + ((u >> 16) & 1u);
  return static_cast<uint16_t>(u >> 16);
}

static float from_bf16(uint16_t h) {
  uint32_t u = static_cast<uint32_t>(h) << 16;
  float f;
  std::memcpy(&f, &u, 4);
  return f;
}

// value rank r contributes to element i in round k
static float contribution(int r, size_t i, int k) { return static_cast<float>((r + 1) * 0.25 + (i % 97) * 0.5 + k); }

static void run_world(int world, size_t elems, int rounds) {
  const size_t arena_bytes = elems * sizeof(float);
  std::vector<char> segment(cb2rt_shm_segment_bytes(world, arena_bytes) + 4096, 0);
  void* base = reinterpret_cast<void*>((reinterpret_cast<uintptr_t>(segment.data()) + 4095) & ~uintptr_t(4095));
  CHECK(cb2rt_shm_format(base, world, arena_bytes) == 0);
  std::vector<int> errors(world, 0);
  std::vector<std::thread> threads;
  for (int r = 0; r < world; ++r) {
    threads.emplace_back([&, r] {
      void* h = cb2rt_shm_attach(base, r, 30.0);
      if (!h) { errors[r] = 1; return; }
      std::mt19937 rng(1234 + r);
      float* mine = static_cast<float*>(cb2rt_shm_arena(h, r));
      uint16_t* mine16 = reinterpret_cast<uint16_t*>(mine);
      for (int k = 0; k < rounds && !errors[r]; ++k) {
        if (rng() % 3 == 0) std::this_thread::sleep_for(std::chrono::microseconds(rng() % 300));
        const int algo = k % 2, dtype = (k / 2) % 2;
        const size_t off = (k % 5) * 3, len = elems - off - (k % 7);
        const float scale = (k % 3 == 0) ? 0.5f : 1.0f;
        for (size_t i = 0; i < len; ++i) {
          if (dtype) mine16[off + i] = to_bf16(contribution(r, i, k));
          else mine[off + i] = contribution(r, i, k);
        }
        float* aux = cb2rt_shm_aux(h);
        aux[0] += 1.0f + r;
        aux[1] += static_cast<float>(k);
        float aux_out[2] = {0, 0};
        if (cb2rt_shm_allreduce(h, algo, dtype, off, len, scale, 2, aux_out) != 0) { errors[r] = 2; break; }
        if (aux_out[0] != world * (world + 1) / 2.0f || aux_out[1] != static_cast<float>(k) * world) errors[r] = 3;
        for (size_t i = 0; i < len && !errors[r]; ++i) {
          float want = 0.f;
          for (int q = 0; q < world; ++q) want += dtype ? bf16_round(contribution(q, i, k)) : contribution(q, i, k);
          want *= scale;
          const float got = dtype ? from_bf16(mine16[off + i]) : mine[off + i];
          if (dtype ? got != bf16_round(want) : got != want) errors[r] = 4;
        }
      }
      // broadcast: rank (world - 1) is the root
      if (!errors[r]) {
        for (size_t i = 0; i < elems; ++i) mine[i] = static_cast<float>(r * 1000 + i);
        if (cb2rt_shm_broadcast(h, world - 1, 16, (elems - 4) * sizeof(float) - 16) != 0) errors[r] = 5;
        const float* f = mine;
        if (r != world - 1 && (f[3] != r * 1000 + 3 || f[4] != (world - 1) * 1000 + 4 ||
                               f[elems - 5] != (world - 1) * 1000 + static_cast<float>(elems - 5) ||
                               f[elems - 4] != r * 1000 + static_cast<float>(elems - 4)))
          errors[r] = 6;
        if (cb2rt_shm_barrier(h) != 0) errors[r] = 7;
      }
      cb2rt_shm_close(h);
    });
  }
  for (auto& t : threads) t.join();
  for (int r = 0; r < world; ++r) {
    if (errors[r]) std::fprintf(stderr, "world %d rank %d: error %d (%s)\n", world, r, errors[r], cb2rt_last_error());
    CHECK(errors[r] == 0);
  }
}

static void missing_rank_times_out() {
  const int world = 3;
  std::vector<char> segment(cb2rt_shm_segment_bytes(world, 1024) + 4096, 0);
  void* base = reinterpret_cast<void*>((reinterpret_cast<uintptr_t>(segment.data()) + 4095) & ~uintptr_t(4095));
  CHECK(cb2rt_shm_format(base, world, 1024) == 0);
  std::vector<int> rcs(world, 0);
  std::vector<std::thread> threads;
  for (int r = 0; r < 2; ++r)                       // rank 2 never shows up
    threads.emplace_back([&, r] {
      void* h = cb2rt_shm_attach(base, r, 0.3);
      float out[1];
      rcs[r] = cb2rt_shm_allreduce(h, 1, 0, 0, 64, 1.0f, 0, out);
      if (rcs[r] == -2 && !std::strstr(cb2rt_last_error(), "for rank 2")) rcs[r] = -99;
      cb2rt_shm_close(h);
    });
  for (auto& t : threads) t.join();
  CHECK(rcs[0] == -2 && rcs[1] == -2);
}

static void argument_checks_and_named_segments() {
  const char* name = "/cb2_test_shm_comm";
  void* a = cb2rt_shm_open(name, 0, 2, 4096, 1, 5.0);
  CHECK(a != nullptr);
  void* b = cb2rt_shm_open(name, 1, 2, 4096, 0, 5.0);
  CHECK(b != nullptr);
  CHECK(cb2rt_shm_open(name, 1, 2, 8192, 0, 5.0) == nullptr);          // different geometry
  CHECK(cb2rt_shm_open("/cb2_test_no_such_segment", 0, 2, 4096, 0, 5.0) == nullptr);
  CHECK(cb2rt_shm_unlink(name) == 0);
  float out[4];
  if (a) {
    CHECK(cb2rt_shm_allreduce(a, 0, 0, 1000, 100, 1.0f, 0, out) == -1);   // past the arena
    CHECK(cb2rt_shm_allreduce(a, 7, 0, 0, 10, 1.0f, 0, out) == -1);
    CHECK(cb2rt_shm_broadcast(a, 5, 0, 16) == -1);
  }
  if (a && b) {                                       // two mappings of the same named segment
    float* fa = static_cast<float*>(cb2rt_shm_arena(a, 0));
    float* fb = static_cast<float*>(cb2rt_shm_arena(b, 1));
    for (int i = 0; i < 8; ++i) { fa[i] = 1.0f + i; fb[i] = 10.0f; }
    std::thread t([&] { CHECK(cb2rt_shm_allreduce(b, 0, 0, 0, 8, 1.0f, 0, out) == 0); });
    float out2[1];
    CHECK(cb2rt_shm_allreduce(a, 0, 0, 0, 8, 1.0f, 0, out2) == 0);
    t.join();
    CHECK(fa[3] == 14.0f && fb[3] == 14.0f);
  }
  cb2rt_shm_close(a);
  cb2rt_shm_close(b);
}

int main() {
  for (int world : {1, 2, 3, 4, 8}) run_world(world, 1 << 12, 24);
  run_world(5, 37, 16);                               // shards that do not divide evenly
  missing_rank_times_out();
  argument_checks_and_named_segments();
  if (g_failed) {
    std::fprintf(stderr, "%d checks failed\n", g_failed);
    return 1;
  }
  std::printf("test_shm_comm: all checks PASSED\n");
  return 0;
}
