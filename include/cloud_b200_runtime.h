/* C ABI of the cloud_b200 C++ runtime library (cloud_b200/_lib/libcloud_b200_rt.so).
 *
 * Four independent parts; each returns 0 / a non-negative count on success unless stated otherwise,
 * and leaves a message for cb2rt_last_error() (thread-local) on failure.
 *
 *   cb2rt_shm_*   host-memory collectives over POSIX shared memory   csrc/runtime/shm_comm.cc
 *   cb2rt_symm_*  CUDA VMM arenas shared between processes, NVLS      csrc/runtime/symm_arena.cc
 *   cb2h5_*       self-contained HDF5 writer / reader                 csrc/runtime/hdf5.cc
 *   cb2mon_*      metrics registry, allow-list, periodic exporter     csrc/runtime/mon_*.cc
 *
 * Reference parity: cb2mon_* is the counterpart of src/cpp/monitoring (stackdriver_client.cc,
 * stackdriver_config.cc, stackdriver_exporter.cc); the other three replace what the reference gets
 * from TensorFlow (collectives), the CUDA driver through TensorFlow / NCCL (peer memory) and
 * h5py / libhdf5 (Keras weight files) and have no source there.
 */
#ifndef CLOUD_B200_RUNTIME_H_
#define CLOUD_B200_RUNTIME_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

const char* cb2rt_last_error(void);

/* ---- shared-memory collectives (CPU ranks of one box) -------------------------------------------
 * A segment holds, per rank, a signal pad (one epoch counter per source rank + 256 fp32 scalars)
 * and an arena of `arena_bytes`.  Rank 0 creates it, the others open it after the creator returned
 * (carry the name over your own control plane), then the name can be unlinked.  Collective calls
 * must be made by every rank in the same order with the same arguments; a rank that does not
 * arrive within the timeout makes the others return -2 (the error names it).  Payloads are
 * addressed in ELEMENTS of the chosen dtype from the start of each rank's arena. */
size_t cb2rt_shm_segment_bytes(int world, size_t arena_bytes);
int cb2rt_shm_max_ranks(void);
int cb2rt_shm_aux_floats(void);
void* cb2rt_shm_open(const char* name, int rank, int world, size_t arena_bytes, int create, double timeout_s);
int cb2rt_shm_unlink(const char* name);
/* Several ranks inside one process on one mapping (tests under ThreadSanitizer): format once, attach per rank. */
int cb2rt_shm_format(void* base, int world, size_t arena_bytes);
void* cb2rt_shm_attach(void* base, int rank, double timeout_s);
void cb2rt_shm_close(void* handle);
int cb2rt_shm_rank(void* handle);
int cb2rt_shm_world(void* handle);
int cb2rt_shm_attached(void* handle);
void* cb2rt_shm_arena(void* handle, int rank);          /* any rank's arena, as mapped here */
float* cb2rt_shm_aux(void* handle);                     /* this rank's scalars */
void cb2rt_shm_set_timeout(void* handle, double seconds);
int cb2rt_shm_barrier(void* handle);
/* algo: 0 one-shot, 1 two-shot.  dtype: 0 fp32, 1 bf16 (fp32 accumulation).  In place; `scale`
 * multiplies the sum.  aux_len > 0: the first aux_len scalars of all pads are summed into aux_out
 * and this rank's are cleared.  len == 0 with aux_len > 0 reduces the scalars only. */
int cb2rt_shm_allreduce(void* handle, int algo, int dtype, size_t off, size_t len, float scale, int aux_len,
                        float* aux_out);
int cb2rt_shm_broadcast(void* handle, int root, size_t byte_off, size_t nbytes);

/* ---- CUDA VMM arenas and NVLS multicast objects (driver API through dlopen'd libcuda) ------------ */
int cb2rt_symm_capabilities(int dev, int* vmm, int* multicast);
int cb2rt_symm_granularity(int dev, int world, size_t* out);
int cb2rt_symm_create(int dev, size_t bytes, uint64_t* handle_out, int* fd_out);
int cb2rt_symm_import(int fd, uint64_t* handle_out);
int cb2rt_symm_map(uint64_t handle, int dev, size_t bytes, size_t align, uint64_t* ptr_out);
int cb2rt_symm_unmap(uint64_t ptr, size_t bytes);
int cb2rt_symm_release(uint64_t handle);
int cb2rt_mc_create(int world, size_t bytes, uint64_t* handle_out, int* fd_out);
int cb2rt_mc_add_device(uint64_t mc, int dev);
int cb2rt_mc_bind(uint64_t mc, uint64_t mem, size_t bytes);
int cb2rt_close_fd(int fd);

/* ---- HDF5 (superblock v0, v1 object headers, symbol-table groups, contiguous datasets) -----------
 * dtype codes: see cloud_b200/io/hdf5.py.  Paths are absolute ("/dense/dense/kernel:0"). */
void* cb2h5_writer_new(void);
void cb2h5_writer_free(void* w);
int cb2h5_writer_add_group(void* w, const char* path);
int cb2h5_writer_add_dataset(void* w, const char* path, int dtype, int rank, const uint64_t* dims, const void* data);
int cb2h5_writer_add_attr(void* w, const char* obj_path, const char* name, int dtype, size_t slen, int rank,
                          const uint64_t* dims, const void* data);
int cb2h5_writer_save(void* w, const char* filename);
void* cb2h5_reader_open(const char* filename);
void cb2h5_reader_close(void* r);
int cb2h5_reader_list(void* r, const char* group_path, char* buf, size_t buflen);
int cb2h5_reader_dataset_info(void* r, const char* path, int* dtype, int* rank, uint64_t* dims /*[8]*/);
int cb2h5_reader_read_dataset(void* r, const char* path, void* out, size_t nbytes);
int cb2h5_reader_attr_names(void* r, const char* obj_path, char* buf, size_t buflen);
int cb2h5_reader_attr_info(void* r, const char* obj_path, const char* name, int* dtype, size_t* slen, int* rank,
                           uint64_t* dims, size_t* nbytes);
int cb2h5_reader_attr_read(void* r, const char* obj_path, const char* name, void* out, size_t nbytes);

/* ---- monitoring ------------------------------------------------------------------------------------
 * Environment: TFC_MONITORING_ENABLED, TFC_MONITORING_PROJECT_ID, TFC_MONITORING_ALLOWLIST,
 * TFC_MONITORING_SINK (and the reference's TF_MONITORING_STACKDRIVER_* spellings). */
void cb2mon_define_counter(const char* name, const char* desc);
void cb2mon_define_gauge(const char* name, const char* desc);
void cb2mon_define_histogram(const char* name, const char* desc, const double* limits, int n);
void cb2mon_counter_inc(const char* name, long long delta);
void cb2mon_gauge_set(const char* name, long long value);
void cb2mon_observe(const char* name, double value);
void cb2mon_reset(void);
int cb2mon_start(void);                 /* 1 if the periodic exporter runs (enabled by the environment) */
void cb2mon_stop(void);
void cb2mon_export_now(void);
int cb2mon_export_count(void);
int cb2mon_is_allowlisted(const char* name);
int cb2mon_config_debug_string(char* buf, size_t n);
int cb2mon_set_sink(int kind /*0 jsonl, 1 prometheus text, 2 memory*/, const char* path);
int cb2mon_memory_sink_dump(int which /*0 time series, 1 descriptors*/, char* buf, size_t n);

#ifdef __cplusplus
}
#endif
#endif  /* CLOUD_B200_RUNTIME_H_ */
