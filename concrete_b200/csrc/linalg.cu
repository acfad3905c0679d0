// linalg.cu — K5: leveled LWE operations the SDFG runtime calls between KS/PBS
// (compilers/concrete-compiler/compiler/lib/Runtime/GPUDFG.cpp:1286,1344,1402,1445).
// Replaces third_party/tfhe-rs/backends/tfhe-cuda-backend/cuda/src/linearalgebra/
// {addition.cuh:16,52; multiplication.cuh:17; negation.cuh:14}; wrapping u64, bit-exact with
// concrete_cpu_{add,add_plaintext,mul_cleartext,negate}_lwe_ciphertext_u64.
#include "common.cuh"
#include "../../include/concrete_b200.h"

namespace {
constexpr int LT = 256;

__global__ void k_negate(u64 *out, const u64 *in, size_t total) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total;
       i += (size_t)gridDim.x * blockDim.x)
    out[i] = (u64)0 - in[i];
}
__global__ void k_add(u64 *out, const u64 *a, const u64 *b,
                      size_t total) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total;
       i += (size_t)gridDim.x * blockDim.x)
    out[i] = a[i] + b[i];
}
// body += plaintext (one plaintext per ciphertext); mask copied
__global__ void k_add_plaintext(u64 *out, const u64 *in,
                                const u64 *pt, u32 lwe_dim, size_t total) {
  const size_t size = (size_t)lwe_dim + 1;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total;
       i += (size_t)gridDim.x * blockDim.x) {
    const size_t ct = i / size, pos = i - ct * size;
    out[i] = in[i] + (pos == lwe_dim ? pt[ct] : (u64)0);
  }
}
// every word of ciphertext c multiplied by cleartext[c]
__global__ void k_mul_cleartext(u64 *out, const u64 *in,
                                const u64 *clear, u32 lwe_dim, size_t total) {
  const size_t size = (size_t)lwe_dim + 1;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total;
       i += (size_t)gridDim.x * blockDim.x)
    out[i] = in[i] * clear[i / size];
}
inline u32 grid_for(size_t total) {
  size_t g = (total + LT - 1) / LT;
  return (u32)(g > 148 * 16 ? 148 * 16 : (g ? g : 1));
}
} // namespace

extern "C" {
void cuda_negate_lwe_ciphertext_vector_64(void *stream, uint32_t gpu_index, void *lwe_array_out,
                                          void *lwe_array_in, uint32_t input_lwe_dimension,
                                          uint32_t input_lwe_ciphertext_count) {
  CB_CHECK(cudaSetDevice(gpu_index));
  const size_t total = (size_t)(input_lwe_dimension + 1) * input_lwe_ciphertext_count;
  if (!total) return;
  k_negate<<<grid_for(total), LT, 0, as_stream(stream)>>>((u64 *)lwe_array_out,
                                                          (const u64 *)lwe_array_in, total);
  CB_CHECK(cudaGetLastError());
  CB_LAUNCHED();
}
void cuda_add_lwe_ciphertext_vector_64(void *stream, uint32_t gpu_index, void *lwe_array_out,
                                       void *lwe_array_in_1, void *lwe_array_in_2,
                                       uint32_t input_lwe_dimension,
                                       uint32_t input_lwe_ciphertext_count) {
  CB_CHECK(cudaSetDevice(gpu_index));
  const size_t total = (size_t)(input_lwe_dimension + 1) * input_lwe_ciphertext_count;
  if (!total) return;
  k_add<<<grid_for(total), LT, 0, as_stream(stream)>>>(
      (u64 *)lwe_array_out, (const u64 *)lwe_array_in_1, (const u64 *)lwe_array_in_2, total);
  CB_CHECK(cudaGetLastError());
  CB_LAUNCHED();
}
void cuda_add_lwe_ciphertext_vector_plaintext_vector_64(void *stream, uint32_t gpu_index,
                                                        void *lwe_array_out, void *lwe_array_in,
                                                        void *plaintext_array_in,
                                                        uint32_t input_lwe_dimension,
                                                        uint32_t input_lwe_ciphertext_count) {
  CB_CHECK(cudaSetDevice(gpu_index));
  const size_t total = (size_t)(input_lwe_dimension + 1) * input_lwe_ciphertext_count;
  if (!total) return;
  k_add_plaintext<<<grid_for(total), LT, 0, as_stream(stream)>>>(
      (u64 *)lwe_array_out, (const u64 *)lwe_array_in, (const u64 *)plaintext_array_in,
      input_lwe_dimension, total);
  CB_CHECK(cudaGetLastError());
  CB_LAUNCHED();
}
void cuda_mult_lwe_ciphertext_vector_cleartext_vector_64(void *stream, uint32_t gpu_index,
                                                         void *lwe_array_out, void *lwe_array_in,
                                                         void *cleartext_array_in,
                                                         uint32_t input_lwe_dimension,
                                                         uint32_t input_lwe_ciphertext_count) {
  CB_CHECK(cudaSetDevice(gpu_index));
  const size_t total = (size_t)(input_lwe_dimension + 1) * input_lwe_ciphertext_count;
  if (!total) return;
  k_mul_cleartext<<<grid_for(total), LT, 0, as_stream(stream)>>>(
      (u64 *)lwe_array_out, (const u64 *)lwe_array_in, (const u64 *)cleartext_array_in,
      input_lwe_dimension, total);
  CB_CHECK(cudaGetLastError());
  CB_LAUNCHED();
}
}
