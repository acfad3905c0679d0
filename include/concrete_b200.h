/* concrete_b200.h — C ABI of the B200-native TFHE hot-path library.
 *
 * Drop-in boundary: every symbol below is what Concrete's compiler runtime
 * (libConcretelangRuntime: lib/Runtime/wrappers.cpp, GPUDFG.cpp, context.h) or the
 * legacy concrete-cuda Rust FFI binds today.  Each declaration cites the reference
 * interface it replaces; paths are relative to /root/reference with
 *   TCB = third_party/tfhe-rs/backends/tfhe-cuda-backend/cuda
 *   RT  = compilers/concrete-compiler/compiler
 *   LEG = backends/concrete-cuda/rust_api/src/cuda_bind.rs
 *
 * Conventions (same as the reference): plain pointers and sizes; `stream` is a
 * cudaStream_t passed as void*; every call starts with cudaSetDevice(gpu_index);
 * all work is asynchronous on `stream`; errors print to stderr and abort()
 * (TCB/include/device.h:14-28).  All ciphertext / index / key pointers are DEVICE
 * pointers unless a parameter says "host".
 */
#ifndef CONCRETE_B200_H
#define CONCRETE_B200_H

#include <stdbool.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ------------------------------------------------------------------ device.h
 * TCB/include/device.h:30-63 (impl TCB/src/device.cu:6-275) */
void *cuda_create_stream(uint32_t gpu_index);
void cuda_destroy_stream(void *stream, uint32_t gpu_index);
void cuda_synchronize_stream(void *stream, uint32_t gpu_index);
void *cuda_malloc(uint64_t size, uint32_t gpu_index);
void *cuda_malloc_async(uint64_t size, void *stream, uint32_t gpu_index);
void cuda_check_valid_malloc(uint64_t size, uint32_t gpu_index);
void cuda_memcpy_async_to_gpu(void *dest, void *src, uint64_t size, void *stream,
                              uint32_t gpu_index);
void cuda_memcpy_async_gpu_to_gpu(void *dest, void *src, uint64_t size, void *stream,
                                  uint32_t gpu_index);
void cuda_memcpy_gpu_to_gpu(void *dest, void *src, uint64_t size, uint32_t gpu_index);
void cuda_memcpy_async_to_cpu(void *dest, const void *src, uint64_t size, void *stream,
                              uint32_t gpu_index);
void cuda_memset_async(void *dest, uint64_t val, uint64_t size, void *stream,
                       uint32_t gpu_index);
int cuda_get_number_of_gpus(void);
void cuda_synchronize_device(uint32_t gpu_index);
void cuda_drop(void *ptr, uint32_t gpu_index);
void cuda_drop_async(void *ptr, void *stream, uint32_t gpu_index);
int cuda_get_max_shared_memory(uint32_t gpu_index); /* LEG:46 ; device.h:66 */

/* --------------------------------------------------------------- keyswitch.h
 * TCB/include/keyswitch.h:14-18 ; called at RT/lib/Runtime/wrappers.cpp:149-151 and
 * RT/lib/Runtime/GPUDFG.cpp:1098-1101.
 * out[lwe_output_indexes[s]] = keyswitch(in[lwe_input_indexes[s]]) for s < num_samples.
 * lwe_array_in  : [*][lwe_dimension_in + 1] u64 ; lwe_array_out : [*][lwe_dimension_out + 1]
 * ksk           : raw tfhe-rs layout [lwe_dimension_in][level_count][lwe_dimension_out + 1],
 *                 memory level index 0 == decomposition level `level_count`.
 * Result is bit-exact with concrete_cpu_keyswitch_lwe_ciphertext_u64. */
void cuda_keyswitch_lwe_ciphertext_vector_64(
    void *stream, uint32_t gpu_index, void *lwe_array_out, void *lwe_output_indexes,
    void *lwe_array_in, void *lwe_input_indexes, void *ksk, uint32_t lwe_dimension_in,
    uint32_t lwe_dimension_out, uint32_t base_log, uint32_t level_count,
    uint32_t num_samples);

/* No counterpart in the reference: added for callers
 * that may not allocate (CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  Same arguments, same device pointers and the same bit-exact result as
 * cuda_keyswitch_lwe_ciphertext_vector_64, computed by the tcgen05 INT8 kernel with its two
 * operand images in a fixed `static __device__` array of
 * CONCRETE_B200_KS_NO_ALLOC_SCRATCH_BYTES (72 MiB, held by every GPU that loads the library
 * until exit) instead of the stream-ordered pool: no cudaMalloc / cudaFree / allocating library
 * routine on this path.
 *   needs   256-rounded(Np * Kp) + 128 * Kp bytes <= the array, Kp = roundup(n_in * level, 128),
 *           Np = roundup(8 * (n_out + 1), 256)  (P2048, n_in 2048 x l 5 -> n_out 742: 61.25 MiB),
 *           and the exact-INT8 range of that kernel: base_log <= 7,
 *           n_in * level * 2^(base_log-1) * 255 < 2^31.
 *   returns 0, or a CONCRETE_B200_NO_ALLOC_* code with a message in error_message (may be NULL);
 *           nothing is launched when it returns non-zero from the argument checks.
 *   sharing every call on a GPU uses the same array.  Each call rewrites every byte it reads, so
 *           stale contents are harmless, but calls that overlap on the GPU (two streams) race:
 *           order them on one stream or with events. */
#define CONCRETE_B200_KS_NO_ALLOC_SCRATCH_BYTES (72ull << 20)
#define CONCRETE_B200_NO_ALLOC_BAD_ARGUMENT 1
#define CONCRETE_B200_NO_ALLOC_UNSUPPORTED 2
#define CONCRETE_B200_NO_ALLOC_TOO_LARGE 3
#define CONCRETE_B200_NO_ALLOC_CUDA_ERROR 4
int concrete_b200_keyswitch_lwe_ciphertext_vector_64_no_alloc(
    void *stream, uint32_t gpu_index, void *lwe_array_out, void *lwe_output_indexes,
    void *lwe_array_in, void *lwe_input_indexes, void *ksk, uint32_t lwe_dimension_in,
    uint32_t lwe_dimension_out, uint32_t base_log, uint32_t level_count,
    uint32_t num_samples, char *error_message, uint64_t error_capacity);

/* ---------------------------------------------------- programmable_bootstrap.h
 * TCB/include/programmable_bootstrap.h:21-24 ; called at RT/include/concretelang/Runtime/
 * context.h:101-105.  `src` is a HOST pointer to the standard-domain BSK
 * [n][l][k+1][k+1][N] u64 ; `dest` is a DEVICE buffer of exactly n*(k+1)^2*l*N*8 bytes
 * whose Fourier layout is private to this library. */
void cuda_convert_lwe_programmable_bootstrap_key_64(
    void *stream, uint32_t gpu_index, void *dest, void *src, uint32_t input_lwe_dim,
    uint32_t glwe_dim, uint32_t level_count, uint32_t polynomial_size);

/* TCB/include/programmable_bootstrap.h:61-64 ; RT/lib/Runtime/wrappers.cpp:234,
 * GPUDFG.cpp:170-188.  Returns an opaque handle in *buffer. */
void scratch_cuda_programmable_bootstrap_64(
    void *stream, uint32_t gpu_index, int8_t **buffer, uint32_t glwe_dimension,
    uint32_t polynomial_size, uint32_t level_count, uint32_t input_lwe_ciphertext_count,
    bool allocate_gpu_memory);

/* TCB/include/programmable_bootstrap.h:74-80 ; RT/lib/Runtime/wrappers.cpp:237-240,
 * GPUDFG.cpp:1214-1218.
 * lut_vector : [num_luts][(k+1)*N] u64 trivial GLWE accumulators ; lut_vector_indexes u64 per
 * sample ; out [*][k*N+1].  lut_count > 1 extracts samples i*lut_stride (i < lut_count) into
 * lwe_array_out + i*num_samples*(k*N+1)  (programmable_bootstrap_classic.cuh:211-259). */
void cuda_programmable_bootstrap_lwe_ciphertext_vector_64(
    void *stream, uint32_t gpu_index, void *lwe_array_out, void *lwe_output_indexes,
    void *lut_vector, void *lut_vector_indexes, void *lwe_array_in,
    void *lwe_input_indexes, void *bootstrapping_key, int8_t *buffer,
    uint32_t lwe_dimension, uint32_t glwe_dimension, uint32_t polynomial_size,
    uint32_t base_log, uint32_t level_count, uint32_t num_samples, uint32_t lut_count,
    uint32_t lut_stride);

/* TCB/include/programmable_bootstrap.h:82-83 */
void cleanup_cuda_programmable_bootstrap(void *stream, uint32_t gpu_index,
                                         int8_t **pbs_buffer);

/* "amortized" entry points, TCB/include/programmable_bootstrap.h:31-54 (the north-star
 * cuda_bootstrap_amortized_* under its current name).  Same kernel, no level_count in scratch. */
void scratch_cuda_programmable_bootstrap_amortized_64(
    void *stream, uint32_t gpu_index, int8_t **pbs_buffer, uint32_t glwe_dimension,
    uint32_t polynomial_size, uint32_t input_lwe_ciphertext_count,
    bool allocate_gpu_memory);
void cuda_programmable_bootstrap_amortized_lwe_ciphertext_vector_64(
    void *stream, uint32_t gpu_index, void *lwe_array_out, void *lwe_output_indexes,
    void *lut_vector, void *lut_vector_indexes, void *lwe_array_in,
    void *lwe_input_indexes, void *bootstrapping_key, int8_t *pbs_buffer,
    uint32_t lwe_dimension, uint32_t glwe_dimension, uint32_t polynomial_size,
    uint32_t base_log, uint32_t level_count, uint32_t num_samples);
void cleanup_cuda_programmable_bootstrap_amortized(void *stream, uint32_t gpu_index,
                                                   int8_t **pbs_buffer);

/* FFT test hook, TCB/include/programmable_bootstrap.h:11-14 (impl bootstrapping_key.cuh:211-324):
 * output[p] = negacyclic product of input1[p] and input2[p] (doubles, N coefficients each). */
void cuda_fourier_polynomial_mul(void *stream, uint32_t gpu_index, void *input1,
                                 void *input2, void *output, uint32_t polynomial_size,
                                 uint32_t total_polynomials);

/* ----------------------------------------------------------- linear_algebra.h
 * TCB/include/linear_algebra.h:15-47 ; called at RT/lib/Runtime/GPUDFG.cpp:1286,1344,1402,1445 */
void cuda_negate_lwe_ciphertext_vector_64(void *stream, uint32_t gpu_index,
                                          void *lwe_array_out, void *lwe_array_in,
                                          uint32_t input_lwe_dimension,
                                          uint32_t input_lwe_ciphertext_count);
void cuda_add_lwe_ciphertext_vector_64(void *stream, uint32_t gpu_index,
                                       void *lwe_array_out, void *lwe_array_in_1,
                                       void *lwe_array_in_2, uint32_t input_lwe_dimension,
                                       uint32_t input_lwe_ciphertext_count);
void cuda_add_lwe_ciphertext_vector_plaintext_vector_64(
    void *stream, uint32_t gpu_index, void *lwe_array_out, void *lwe_array_in,
    void *plaintext_array_in, uint32_t input_lwe_dimension,
    uint32_t input_lwe_ciphertext_count);
void cuda_mult_lwe_ciphertext_vector_cleartext_vector_64(
    void *stream, uint32_t gpu_index, void *lwe_array_out, void *lwe_array_in,
    void *cleartext_array_in, uint32_t input_lwe_dimension,
    uint32_t input_lwe_ciphertext_count);

/* ------------------------------------------------- legacy concrete-cuda names
 * LEG:53-62,106-116,187-205,209-213,320-331.  The 10-argument legacy keyswitch collides with
 * the current 12-argument symbol of the same name, so it is exported with a `_legacy` suffix
 * (SURVEY.md §0.2). */
void cuda_convert_lwe_bootstrap_key_64(void *dest, void *src, void *v_stream,
                                       uint32_t gpu_index, uint32_t input_lwe_dim,
                                       uint32_t glwe_dim, uint32_t level_count,
                                       uint32_t polynomial_size);
void scratch_cuda_bootstrap_low_latency_64(void *v_stream, uint32_t gpu_index,
                                           int8_t **pbs_buffer, uint32_t glwe_dimension,
                                           uint32_t polynomial_size, uint32_t level_count,
                                           uint32_t input_lwe_ciphertext_count,
                                           uint32_t max_shared_memory,
                                           bool allocate_gpu_memory);
void cuda_bootstrap_low_latency_lwe_ciphertext_vector_64(
    void *v_stream, uint32_t gpu_index, void *lwe_array_out, void *lut_vector,
    void *lut_vector_indexes, void *lwe_array_in, void *bootstrapping_key,
    int8_t *pbs_buffer, uint32_t lwe_dimension, uint32_t glwe_dimension,
    uint32_t polynomial_size, uint32_t base_log, uint32_t level, uint32_t num_samples,
    uint32_t num_lut_vectors, uint32_t lwe_idx, uint32_t max_shared_memory);
void cleanup_cuda_bootstrap_low_latency(void *v_stream, uint32_t gpu_index,
                                        int8_t **pbs_buffer);
void scratch_cuda_bootstrap_amortized_64(void *v_stream, uint32_t gpu_index,
                                         int8_t **pbs_buffer, uint32_t glwe_dimension,
                                         uint32_t polynomial_size,
                                         uint32_t input_lwe_ciphertext_count,
                                         uint32_t max_shared_memory,
                                         bool allocate_gpu_memory);
void cuda_bootstrap_amortized_lwe_ciphertext_vector_64(
    void *v_stream, uint32_t gpu_index, void *lwe_array_out, void *lut_vector,
    void *lut_vector_indexes, void *lwe_array_in, void *bootstrapping_key,
    int8_t *pbs_buffer, uint32_t lwe_dimension, uint32_t glwe_dimension,
    uint32_t polynomial_size, uint32_t base_log, uint32_t level, uint32_t num_samples,
    uint32_t num_lut_vectors, uint32_t lwe_idx, uint32_t max_shared_memory);
void cleanup_cuda_bootstrap_amortized(void *v_stream, uint32_t gpu_index,
                                      int8_t **pbs_buffer);
void cuda_keyswitch_lwe_ciphertext_vector_64_legacy(
    void *v_stream, uint32_t gpu_index, void *lwe_array_out, void *lwe_array_in,
    void *keyswitch_key, uint32_t input_lwe_dimension, uint32_t output_lwe_dimension,
    uint32_t base_log, uint32_t level_count, uint32_t num_samples);

/* ----------------------------------------------------- fused chain (extension)
 * KS -> PBS chained on device with no host round trip — what GPUDFG.cpp:1057-1250 does per
 * chunk with two library calls and an intermediate buffer.  `ks_tmp` is a device scratch of
 * num_samples*(lwe_dimension+1) u64. */
void cuda_keyswitch_bootstrap_lwe_ciphertext_vector_64(
    void *stream, uint32_t gpu_index, void *lwe_array_out, void *lut_vector,
    void *lut_vector_indexes, void *lwe_array_in, void *ksk, void *bootstrapping_key,
    int8_t *pbs_buffer, void *ks_tmp, uint32_t big_lwe_dimension, uint32_t lwe_dimension,
    uint32_t ks_base_log, uint32_t ks_level_count, uint32_t glwe_dimension,
    uint32_t polynomial_size, uint32_t base_log, uint32_t level_count,
    uint32_t num_samples);

/* Library self-description (not in the reference): build tag + last-launch kernel counter,
 * used by bench.py to report `gpu_launches`. */
const char *concrete_b200_version(void);
/* ---- seeded (compressed) keys, decompressed on the device ----------------
 * Replace concrete_cpu_decompress_seeded_lwe_bootstrap_key_u64
 * (backends/concrete-cpu/implementation/src/c_api/bootstrap.rs:169-237) and
 * concrete_cpu_decompress_seeded_lwe_keyswitch_key_u64 (c_api/keyswitch.rs) + the full-key upload
 * of RuntimeContext::get_{bsk,ksk}_gpu (context.h:86-145): the caller passes the seeded key (bodies
 * only: n*l*(k+1)*N words, resp. n_in*l words; host or device pointer) and the 128-bit
 * compression seed (Uint128 little_endian_bytes as two u64 halves); the masks are regenerated on
 * the device with the reference's AES-128-CTR generator (concrete-csprng).  `dest` is a device
 * buffer of the full key size; the _convert_ form leaves the private Fourier layout in it, as
 * cuda_convert_lwe_programmable_bootstrap_key_64 does. */
void cuda_decompress_seeded_lwe_bootstrap_key_64(void *stream, uint32_t gpu_index, void *dest,
                                                 const void *seeded_bsk, uint32_t input_lwe_dim,
                                                 uint32_t glwe_dim, uint32_t level_count,
                                                 uint32_t polynomial_size, uint64_t seed_lo,
                                                 uint64_t seed_hi);
void cuda_decompress_seeded_lwe_keyswitch_key_64(void *stream, uint32_t gpu_index, void *dest,
                                                 const void *seeded_ksk, uint32_t lwe_dim_in,
                                                 uint32_t lwe_dim_out, uint32_t level_count,
                                                 uint64_t seed_lo, uint64_t seed_hi);
void cuda_convert_seeded_lwe_programmable_bootstrap_key_64(
    void *stream, uint32_t gpu_index, void *dest, const void *seeded_bsk, uint32_t input_lwe_dim,
    uint32_t glwe_dim, uint32_t level_count, uint32_t polynomial_size, uint64_t seed_lo,
    uint64_t seed_hi);

/* ---- WoP-PBS on the device ---------------------------------------------------
 * The reference runs this path on the CPU only (memref_wop_pbs_crt_buffer,
 * compilers/concrete-compiler/compiler/lib/Runtime/wrappers.cpp:855-997, over
 * concrete_cpu_extract_bit_lwe_ciphertext_u64 and
 * concrete_cpu_circuit_bootstrap_boolean_vertical_packing_lwe_ciphertext_u64,
 * backends/concrete-cpu/implementation/src/c_api/wop_pbs.rs:103-412).  Same arithmetic, same
 * argument meaning; all array arguments are DEVICE pointers except the small parameter arrays
 * (`number_of_bits`, `crt_decomposition`: host).  `fourier_bsk` is the key produced by
 * cuda_convert_lwe_programmable_bootstrap_key_64, `ksk` the raw big->small keyswitching key,
 * `pfpksk` the (k+1) private functional packing keyswitching keys of the circuit bootstrap, raw
 * tfhe-rs layout [k+1][kN+1][level][(k+1)N].
 *   extract_bits: block i yields number_of_bits[i] ciphertexts (MSB first) with
 *     delta_log = 64 - number_of_bits[i]; the bit list is laid out as the runtime expects, the LAST
 *     block first (wrappers.cpp:903-909); the input is used as is (the caller applies the rounding
 *     trick of wrappers.cpp:928-931 if it wants it — cuda_wop_pbs_crt_... does);
 *   circuit_bootstrap_vertical_packing: `bits_in` [number_of_bits][n+1], `luts`
 *     [lut_count][2^number_of_bits], out [lut_count][kN+1];
 *   wop_pbs_crt: the whole memref_wop_pbs_crt_buffer, in/out [blocks][kN+1],
 *     lut_ct [blocks][2^total_bits]. */
void cuda_extract_bits_lwe_ciphertext_vector_64(
    void *stream, uint32_t gpu_index, void *bits_out, const void *lwe_array_in,
    const uint32_t *number_of_bits, uint32_t blocks, const void *fourier_bsk, const void *ksk,
    uint32_t lwe_small_dim, uint32_t glwe_dim, uint32_t polynomial_size, uint32_t bsk_level,
    uint32_t bsk_base_log, uint32_t ksk_level, uint32_t ksk_base_log);
void cuda_circuit_bootstrap_vertical_packing_lwe_ciphertext_vector_64(
    void *stream, uint32_t gpu_index, void *lwe_array_out, const void *bits_in, const void *luts,
    const void *fourier_bsk, const void *pfpksk, uint32_t number_of_bits, uint32_t lut_count,
    uint64_t lut_size, uint32_t lwe_small_dim, uint32_t glwe_dim, uint32_t polynomial_size,
    uint32_t bsk_level, uint32_t bsk_base_log, uint32_t pksk_level, uint32_t pksk_base_log,
    uint32_t cbs_level, uint32_t cbs_base_log);
void cuda_wop_pbs_crt_lwe_ciphertext_vector_64(
    void *stream, uint32_t gpu_index, void *lwe_array_out, const void *lwe_array_in,
    const void *lut_ct, const uint64_t *crt_decomposition, uint32_t blocks, const void *fourier_bsk,
    const void *ksk, const void *pfpksk, uint32_t lwe_small_dim, uint32_t glwe_dim,
    uint32_t polynomial_size, uint32_t bsk_level, uint32_t bsk_base_log, uint32_t ksk_level,
    uint32_t ksk_base_log, uint32_t pksk_level, uint32_t pksk_base_log, uint32_t cbs_level,
    uint32_t cbs_base_log);

uint64_t concrete_b200_kernel_launches(void);
/* TEST hooks (not in the reference): pin the kernel family so that the tests keep every kernel
 * covered.  PBS: 0 automatic, 1 first generic kernel, 2 group kernel, 3 cluster kernel, 4 v4
 * (pbs_v4.cuh) wherever it has the shape; must be set BEFORE the bootstrapping key is converted
 * (the Fourier key layout belongs to the family).  Keyswitch: 0 automatic, 1 integer pipe, 2 INT8
 * tensor cores (tcgen05.mma kind::i8, TMEM accumulators) whenever exact. */
void concrete_b200_set_pbs_plan(int plan);
void concrete_b200_set_ks_plan(int plan);
/* TUNING hook: alternative instantiations of v4 kernels, 0 = the default.  N = 16384: 2 = 16 points
 * per thread (the default has 32), 1 = the same with packed decomposition states.  k = 2, N = 1024:
 * 3 = 8 points per thread and 3 ciphertexts per CTA.  All of them change the Fourier key layout, so
 * set the hook BEFORE the key is converted (and with plan 4 for k = 2: the automatic plan's
 * small-batch kernel reads the default layout only). */
void concrete_b200_set_v4_alt(int alt);
/* FP64 FMA peak probe: `blocks` CTAs x 256 threads x iters x 16 DFMA (32 flop) per thread;
 * bench.py times it with CUDA events for the roofline denominator. */
void concrete_b200_fp64_peak(void *stream, uint32_t gpu_index, void *scratch, uint32_t blocks,
                             uint32_t iters);

#ifdef __cplusplus
}
#endif
#endif /* CONCRETE_B200_H */
