// Unit test of the self-contained HDF5 writer/reader (hdf5.cc) through its C ABI: a Keras-shaped
// weight file round trip, and robustness of the reader against truncated and bit-flipped files
// (it must fail cleanly, never read out of bounds - the sanitizer builds of this test check that).
// New tier: the reference leaves HDF5 to h5py/libhdf5.
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <unistd.h>
#include <fstream>
#include <string>
#include <vector>

extern "C" {
const char* cb2rt_last_error();
void* cb2h5_writer_new();
void cb2h5_writer_free(void* w);
int cb2h5_writer_add_group(void* w, const char* path);
int cb2h5_writer_add_dataset(void* w, const char* path, int dtype, int rank, const uint64_t* dims,
                             const void* data);
int cb2h5_writer_add_attr(void* w, const char* obj_path, const char* name, int dtype, size_t slen,
                          int rank, const uint64_t* dims, const void* data);
int cb2h5_writer_save(void* w, const char* filename);
void* cb2h5_reader_open(const char* filename);
void cb2h5_reader_close(void* r);
int cb2h5_reader_list(void* r, const char* group_path, char* buf, size_t buflen);
int cb2h5_reader_dataset_info(void* r, const char* path, int* dtype, int* rank, uint64_t* dims);
int cb2h5_reader_read_dataset(void* r, const char* path, void* out, size_t nbytes);
int cb2h5_reader_attr_names(void* r, const char* obj_path, char* buf, size_t buflen);
int cb2h5_reader_attr_info(void* r, const char* obj_path, const char* name, int* dtype, size_t* slen,
                           int* rank, uint64_t* dims, size_t* nbytes);
int cb2h5_reader_attr_read(void* r, const char* obj_path, const char* name, void* out, size_t nbytes);
}

static int g_failed = 0;
#define CHECK(cond)                                                              \
  do {                                                                           \
    if (!(cond)) {                                                               \
      std::fprintf(stderr, "CHECK failed %s:%d: %s (last error: %s)\n", __FILE__, __LINE__, #cond, \
                   cb2rt_last_error());                                          \
      ++g_failed;                                                                \
    }                                                                            \
  } while (0)

enum { DT_F32 = 0, DT_F64 = 1, DT_I32 = 2, DT_I64 = 3, DT_U8 = 4, DT_STR = 5 };

static std::string tmp_path(const char* name) {
  const char* dir = std::getenv("TMPDIR");
  return std::string(dir ? dir : "/tmp") + "/cb2_h5_" + std::to_string(static_cast<long>(getpid())) + "_" + name;
}

static std::vector<float> kernel_values(size_t n) {
  std::vector<float> v(n);
  for (size_t i = 0; i < n; ++i) v[i] = 0.25f * static_cast<float>(i) - 3.f;
  return v;
}

static void write_weights_file(const std::string& path) {
  void* w = cb2h5_writer_new();
  CHECK(cb2h5_writer_add_group(w, "/dense") == 0);
  CHECK(cb2h5_writer_add_group(w, "/dense/dense") == 0);
  const uint64_t kd[2] = {7, 5};
  const std::vector<float> k = kernel_values(35);
  CHECK(cb2h5_writer_add_dataset(w, "/dense/dense/kernel:0", DT_F32, 2, kd, k.data()) == 0);
  const uint64_t bd[1] = {5};
  const double bias[5] = {1, 2, 3, 4, 5};
  CHECK(cb2h5_writer_add_dataset(w, "/dense/dense/bias:0", DT_F64, 1, bd, bias) == 0);
  const int64_t step = 1234567890123LL;
  CHECK(cb2h5_writer_add_dataset(w, "/step", DT_I64, 0, nullptr, &step) == 0);
  // root attributes in the Keras layout: fixed-length string array + scalar string + number
  char names[2][8] = {"dense", "bn"};
  const uint64_t nd[1] = {2};
  CHECK(cb2h5_writer_add_attr(w, "/", "layer_names", DT_STR, 8, 1, nd, names) == 0);
  const char backend[] = "cloud_b200";
  CHECK(cb2h5_writer_add_attr(w, "/", "backend", DT_STR, sizeof(backend), 0, nullptr, backend) == 0);
  const int32_t version = 3;
  CHECK(cb2h5_writer_add_attr(w, "/", "version", DT_I32, 0, 0, nullptr, &version) == 0);
  char wn[2][24] = {"dense/kernel:0", "dense/bias:0"};
  CHECK(cb2h5_writer_add_attr(w, "/dense", "weight_names", DT_STR, 24, 1, nd, wn) == 0);
  // replacing an attribute keeps one copy
  const int32_t version2 = 4;
  CHECK(cb2h5_writer_add_attr(w, "/", "version", DT_I32, 0, 0, nullptr, &version2) == 0);
  // an attribute that cannot fit an object-header message is refused, not truncated
  std::vector<char> big(70000, 'x');
  CHECK(cb2h5_writer_add_attr(w, "/", "model_config", DT_STR, big.size(), 0, nullptr, big.data()) == 2);
  CHECK(cb2h5_writer_add_attr(w, "/nope", "a", DT_I32, 0, 0, nullptr, &version) != 0);
  CHECK(cb2h5_writer_save(w, path.c_str()) == 0);
  cb2h5_writer_free(w);
}

static void test_round_trip() {
  const std::string path = tmp_path("rt.h5");
  write_weights_file(path);
  std::ifstream f(path, std::ios::binary);
  char sig[8];
  f.read(sig, 8);
  CHECK(std::memcmp(sig, "\x89HDF\r\n\x1a\n", 8) == 0);

  void* r = cb2h5_reader_open(path.c_str());
  CHECK(r != nullptr);
  if (!r) return;
  char buf[512];
  CHECK(cb2h5_reader_list(r, "/", buf, sizeof(buf)) >= 0);
  const std::string root(buf);
  CHECK(root.find("dense") != std::string::npos && root.find("step") != std::string::npos);
  int dtype = -1, rank = -1;
  uint64_t dims[8] = {0};
  CHECK(cb2h5_reader_dataset_info(r, "/dense/dense/kernel:0", &dtype, &rank, dims) == 0);
  CHECK(dtype == DT_F32 && rank == 2 && dims[0] == 7 && dims[1] == 5);
  std::vector<float> got(35);
  CHECK(cb2h5_reader_read_dataset(r, "/dense/dense/kernel:0", got.data(), 35 * sizeof(float)) == 0);
  CHECK(got == kernel_values(35));
  CHECK(cb2h5_reader_read_dataset(r, "/dense/dense/kernel:0", got.data(), 34 * sizeof(float)) != 0);
  double bias[5];
  CHECK(cb2h5_reader_read_dataset(r, "/dense/dense/bias:0", bias, sizeof(bias)) == 0 && bias[4] == 5.0);
  int64_t step = 0;
  CHECK(cb2h5_reader_dataset_info(r, "/step", &dtype, &rank, dims) == 0 && rank == 0 && dtype == DT_I64);
  CHECK(cb2h5_reader_read_dataset(r, "/step", &step, sizeof(step)) == 0 && step == 1234567890123LL);
  CHECK(cb2h5_reader_dataset_info(r, "/missing", &dtype, &rank, dims) != 0);

  CHECK(cb2h5_reader_attr_names(r, "/", buf, sizeof(buf)) >= 0);
  const std::string an(buf);
  CHECK(an.find("layer_names") != std::string::npos && an.find("backend") != std::string::npos);
  size_t slen = 0, nbytes = 0;
  CHECK(cb2h5_reader_attr_info(r, "/", "layer_names", &dtype, &slen, &rank, dims, &nbytes) == 0);
  CHECK(dtype == DT_STR && slen == 8 && rank == 1 && dims[0] == 2 && nbytes == 16);
  char names[16];
  CHECK(cb2h5_reader_attr_read(r, "/", "layer_names", names, 16) == 0);
  CHECK(std::strcmp(names, "dense") == 0 && std::strcmp(names + 8, "bn") == 0);
  int32_t version = 0;
  CHECK(cb2h5_reader_attr_info(r, "/", "version", &dtype, &slen, &rank, dims, &nbytes) == 0 && nbytes == 4);
  CHECK(cb2h5_reader_attr_read(r, "/", "version", &version, 4) == 0 && version == 4);
  CHECK(cb2h5_reader_attr_info(r, "/dense", "weight_names", &dtype, &slen, &rank, dims, &nbytes) == 0);
  CHECK(slen == 24 && dims[0] == 2);
  CHECK(cb2h5_reader_attr_info(r, "/", "model_config", &dtype, &slen, &rank, dims, &nbytes) != 0);
  cb2h5_reader_close(r);
  std::remove(path.c_str());
}

// The reader is handed broken files: every truncation length in steps, and single-byte corruptions
// across the metadata.  Opening may fail or succeed; walking whatever opened must not crash.
static void walk(void* r) {
  char buf[256];
  int dtype, rank;
  uint64_t dims[8];
  size_t slen, nbytes;
  if (cb2h5_reader_list(r, "/", buf, sizeof(buf)) < 0) return;
  cb2h5_reader_attr_names(r, "/", buf, sizeof(buf));
  cb2h5_reader_attr_info(r, "/", "layer_names", &dtype, &slen, &rank, dims, &nbytes);
  cb2h5_reader_list(r, "/dense", buf, sizeof(buf));
  cb2h5_reader_list(r, "/dense/dense", buf, sizeof(buf));
  if (cb2h5_reader_dataset_info(r, "/dense/dense/kernel:0", &dtype, &rank, dims) == 0) {
    std::vector<uint8_t> out(35 * 4);
    cb2h5_reader_read_dataset(r, "/dense/dense/kernel:0", out.data(), out.size());
  }
}

static void test_broken_files_fail_cleanly() {
  const std::string good = tmp_path("good.h5");
  write_weights_file(good);
  std::ifstream f(good, std::ios::binary);
  std::vector<char> bytes((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>());
  CHECK(bytes.size() > 512);
  const std::string bad = tmp_path("bad.h5");
  int opened = 0, refused = 0;
  auto try_file = [&](const std::vector<char>& data) {
    std::ofstream o(bad, std::ios::binary | std::ios::trunc);
    o.write(data.data(), static_cast<std::streamsize>(data.size()));
    o.close();
    void* r = cb2h5_reader_open(bad.c_str());
    if (r) {
      ++opened;
      walk(r);
      cb2h5_reader_close(r);
    } else {
      ++refused;
    }
  };
  for (size_t len = 0; len < bytes.size(); len += 37)
    try_file(std::vector<char>(bytes.begin(), bytes.begin() + static_cast<long>(len)));
  uint32_t rng = 12345;
  const size_t meta = bytes.size() < 4096 ? bytes.size() : 4096;
  for (int i = 0; i < 400; ++i) {
    rng = rng * 1664525u + 1013904223u;
    std::vector<char> c = bytes;
    c[(rng >> 8) % meta] = static_cast<char>(rng & 0xff);
    try_file(c);
  }
  CHECK(refused > 0);
  CHECK(cb2h5_reader_open("/nonexistent/dir/file.h5") == nullptr);
  std::printf("broken files: %d refused, %d opened and walked\n", refused, opened);
  std::remove(bad.c_str());
  std::remove(good.c_str());
}

int main() {
  test_round_trip();
  test_broken_files_fail_cleanly();
  if (g_failed) {
    std::fprintf(stderr, "%d HDF5 CHECKS FAILED\n", g_failed);
    return 1;
  }
  std::printf("ALL HDF5 TESTS PASSED\n");
  return 0;
}
