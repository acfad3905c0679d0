// legacy.cu — legacy concrete-cuda entry points (north-star names) + the fused KS->PBS chain.
// Signatures: backends/concrete-cuda/rust_api/src/cuda_bind.rs:53-62,106-116,187-205,209-213,
// 320-331.  They are thin adapters over the current-generation entry points: identity index
// arrays, `lwe_idx` as an offset into lut_vector_indexes, `max_shared_memory` ignored.
#include "common.cuh"
#include "../../include/concrete_b200.h"
#include <map>
#include <mutex>
#include <vector>

void cb_keyswitch_launch(cudaStream_t stream, u64 *out, const u64 *out_idx, const u64 *in,
                         const u64 *in_idx, const u64 *ksk, u32 n_in, u32 n_out, u32 base_log,
                         u32 level, u32 num_samples);
void cb_pbs_scratch(cudaStream_t stream, int gpu, int8_t **buffer, u32 k, u32 N, u32 level,
                    u32 count, bool allocate);
void cb_pbs_cleanup(cudaStream_t stream, int gpu, int8_t **buffer);
void cb_pbs_launch(cudaStream_t stream, int gpu, u64 *lwe_out, const u64 *out_idx,
                   const u64 *lut_vector, const u64 *lut_idx, const u64 *lwe_in,
                   const u64 *in_idx, const double2 *bsk, int8_t *buffer, u32 n, u32 k, u32 N,
                   u32 base_log, u32 level, u32 num_samples, u32 lut_count, u32 lut_stride);
void cb_bsk_convert(cudaStream_t stream, int gpu, void *dest, const void *src_host, u32 n, u32 k,
                    u32 level, u32 N, bool src_on_device);

namespace {
__global__ void k_iota(u64 *p, u32 n) {
  for (u32 i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) p[i] = i;
}

// Per-device identity index array 0..cap-1.  The legacy ABI has no index arguments; the current
// kernels take them.  Arrays are filled on a private non-blocking stream that only the calling
// host thread waits for (no device-wide synchronisation, nothing on the NULL stream), and an
// array that has been handed out is never freed: growth allocates a larger one and keeps the old
// pointers valid for callers that fetched them concurrently (bounded: capacities double).
std::mutex g_iota_mutex;
struct Iota { u64 *ptr = nullptr; u32 cap = 0; cudaStream_t fill = nullptr; std::vector<u64 *> old; };
std::map<int, Iota> g_iota;

const u64 *identity_indexes(int gpu, u32 n) {
  std::lock_guard<std::mutex> g(g_iota_mutex);
  Iota &it = g_iota[gpu];
  if (it.cap >= n) return it.ptr;
  u32 cap = 1u << 20; // 8 MB: covers every batch the runtime sends without ever growing
  while (cap < n) cap <<= 1;
  if (!it.fill) CB_CHECK(cudaStreamCreateWithFlags(&it.fill, cudaStreamNonBlocking));
  if (it.ptr) it.old.push_back(it.ptr);
  CB_CHECK(cudaMalloc((void **)&it.ptr, (size_t)cap * sizeof(u64)));
  k_iota<<<(cap + 255) / 256, 256, 0, it.fill>>>(it.ptr, cap);
  CB_CHECK(cudaGetLastError());
  CB_CHECK(cudaStreamSynchronize(it.fill));
  CB_LAUNCHED();
  it.cap = cap;
  return it.ptr;
}
} // namespace

const u64 *cb_identity_indexes(int gpu, u32 n) { return identity_indexes(gpu, n); }

extern "C" {

void cuda_convert_lwe_bootstrap_key_64(void *dest, void *src, void *v_stream, uint32_t gpu_index,
                                       uint32_t input_lwe_dim, uint32_t glwe_dim,
                                       uint32_t level_count, uint32_t polynomial_size) {
  cuda_convert_lwe_programmable_bootstrap_key_64(v_stream, gpu_index, dest, src, input_lwe_dim,
                                                 glwe_dim, level_count, polynomial_size);
}

void scratch_cuda_bootstrap_low_latency_64(void *v_stream, uint32_t gpu_index, int8_t **pbs_buffer,
                                           uint32_t glwe_dimension, uint32_t polynomial_size,
                                           uint32_t level_count,
                                           uint32_t input_lwe_ciphertext_count,
                                           uint32_t max_shared_memory, bool allocate_gpu_memory) {
  (void)max_shared_memory;
  scratch_cuda_programmable_bootstrap_64(v_stream, gpu_index, pbs_buffer, glwe_dimension,
                                         polynomial_size, level_count,
                                         input_lwe_ciphertext_count, allocate_gpu_memory);
}

void cuda_bootstrap_low_latency_lwe_ciphertext_vector_64(
    void *v_stream, uint32_t gpu_index, void *lwe_array_out, void *lut_vector,
    void *lut_vector_indexes, void *lwe_array_in, void *bootstrapping_key, int8_t *pbs_buffer,
    uint32_t lwe_dimension, uint32_t glwe_dimension, uint32_t polynomial_size, uint32_t base_log,
    uint32_t level, uint32_t num_samples, uint32_t num_lut_vectors, uint32_t lwe_idx,
    uint32_t max_shared_memory) {
  (void)max_shared_memory;
  (void)num_lut_vectors;
  CB_CHECK(cudaSetDevice(gpu_index));
  const u64 *iota = identity_indexes(gpu_index, num_samples);
  cb_pbs_launch(as_stream(v_stream), gpu_index, (u64 *)lwe_array_out, iota,
                (const u64 *)lut_vector, (const u64 *)lut_vector_indexes + lwe_idx,
                (const u64 *)lwe_array_in, iota, (const double2 *)bootstrapping_key, pbs_buffer,
                lwe_dimension, glwe_dimension, polynomial_size, base_log, level, num_samples, 1,
                1);
}

void cleanup_cuda_bootstrap_low_latency(void *v_stream, uint32_t gpu_index, int8_t **pbs_buffer) {
  cleanup_cuda_programmable_bootstrap(v_stream, gpu_index, pbs_buffer);
}

void scratch_cuda_bootstrap_amortized_64(void *v_stream, uint32_t gpu_index, int8_t **pbs_buffer,
                                         uint32_t glwe_dimension, uint32_t polynomial_size,
                                         uint32_t input_lwe_ciphertext_count,
                                         uint32_t max_shared_memory, bool allocate_gpu_memory) {
  (void)max_shared_memory;
  scratch_cuda_programmable_bootstrap_amortized_64(v_stream, gpu_index, pbs_buffer,
                                                   glwe_dimension, polynomial_size,
                                                   input_lwe_ciphertext_count,
                                                   allocate_gpu_memory);
}

void cuda_bootstrap_amortized_lwe_ciphertext_vector_64(
    void *v_stream, uint32_t gpu_index, void *lwe_array_out, void *lut_vector,
    void *lut_vector_indexes, void *lwe_array_in, void *bootstrapping_key, int8_t *pbs_buffer,
    uint32_t lwe_dimension, uint32_t glwe_dimension, uint32_t polynomial_size, uint32_t base_log,
    uint32_t level, uint32_t num_samples, uint32_t num_lut_vectors, uint32_t lwe_idx,
    uint32_t max_shared_memory) {
  cuda_bootstrap_low_latency_lwe_ciphertext_vector_64(
      v_stream, gpu_index, lwe_array_out, lut_vector, lut_vector_indexes, lwe_array_in,
      bootstrapping_key, pbs_buffer, lwe_dimension, glwe_dimension, polynomial_size, base_log,
      level, num_samples, num_lut_vectors, lwe_idx, max_shared_memory);
}

void cleanup_cuda_bootstrap_amortized(void *v_stream, uint32_t gpu_index, int8_t **pbs_buffer) {
  cleanup_cuda_programmable_bootstrap(v_stream, gpu_index, pbs_buffer);
}

void cuda_keyswitch_lwe_ciphertext_vector_64_legacy(void *v_stream, uint32_t gpu_index,
                                                    void *lwe_array_out, void *lwe_array_in,
                                                    void *keyswitch_key,
                                                    uint32_t input_lwe_dimension,
                                                    uint32_t output_lwe_dimension,
                                                    uint32_t base_log, uint32_t level_count,
                                                    uint32_t num_samples) {
  CB_CHECK(cudaSetDevice(gpu_index));
  const u64 *iota = identity_indexes(gpu_index, num_samples);
  cb_keyswitch_launch(as_stream(v_stream), (u64 *)lwe_array_out, iota, (const u64 *)lwe_array_in,
                      iota, (const u64 *)keyswitch_key, input_lwe_dimension,
                      output_lwe_dimension, base_log, level_count, num_samples);
}

// KS -> PBS on device, intermediate small-key ciphertexts never leave HBM.
void cuda_keyswitch_bootstrap_lwe_ciphertext_vector_64(
    void *stream, uint32_t gpu_index, void *lwe_array_out, void *lut_vector,
    void *lut_vector_indexes, void *lwe_array_in, void *ksk, void *bootstrapping_key,
    int8_t *pbs_buffer, void *ks_tmp, uint32_t big_lwe_dimension, uint32_t lwe_dimension,
    uint32_t ks_base_log, uint32_t ks_level_count, uint32_t glwe_dimension,
    uint32_t polynomial_size, uint32_t base_log, uint32_t level_count, uint32_t num_samples) {
  CB_CHECK(cudaSetDevice(gpu_index));
  const u64 *iota = identity_indexes(gpu_index, num_samples);
  cb_keyswitch_launch(as_stream(stream), (u64 *)ks_tmp, iota, (const u64 *)lwe_array_in, iota,
                      (const u64 *)ksk, big_lwe_dimension, lwe_dimension, ks_base_log,
                      ks_level_count, num_samples);
  cb_pbs_launch(as_stream(stream), gpu_index, (u64 *)lwe_array_out, iota,
                (const u64 *)lut_vector, (const u64 *)lut_vector_indexes, (const u64 *)ks_tmp,
                iota, (const double2 *)bootstrapping_key, pbs_buffer, lwe_dimension,
                glwe_dimension, polynomial_size, base_log, level_count, num_samples, 1, 1);
}

} // extern "C"
