/* concrete_b200_runtime.h — runtime wrapper layer (MLIR memref ABI) of the drop-in boundary.
 *
 * These are the functions Concrete's generated code calls (declared in the reference at
 * compilers/concrete-compiler/compiler/include/concretelang/Runtime/wrappers.h and defined at
 * lib/Runtime/wrappers.cpp:70-363,745-849).  Prototypes are identical; `RuntimeContext` is the
 * runtime's context class (include/concretelang/Runtime/context.h:42-154) — built here against
 * the shim in concrete_b200/csrc/runtime_context.h.  2-D memrefs are passed as
 * (allocated, aligned, offset, size0, size1, stride0, stride1), 1-D as (.., size, stride);
 * all memref pointers are HOST pointers.
 */
#ifndef CONCRETE_B200_RUNTIME_H
#define CONCRETE_B200_RUNTIME_H
#include <stdint.h>

#ifdef __cplusplus
namespace mlir { namespace concretelang { struct RuntimeContext; } }
typedef mlir::concretelang::RuntimeContext cb_runtime_context;
extern "C" {
#else
typedef struct RuntimeContext cb_runtime_context;
#endif

/* context factory: stands in for RuntimeContext(ServerKeyset) (lib/Runtime/context.cpp:32-57).
 * Key buffers are HOST pointers in the reference's raw layouts and stay owned by the caller. */
void *concrete_b200_context_create(void);
void concrete_b200_context_destroy(void *ctx);
uint64_t concrete_b200_context_add_ksk(void *ctx, const uint64_t *data, uint64_t len_words);
uint64_t concrete_b200_context_add_bsk(void *ctx, const uint64_t *data, uint64_t len_words);
/* seeded (compressed) keys: bodies + 128-bit compression seed; `full_len_words` is the word count
 * of the decompressed key.  The masks are regenerated on every device (concrete_b200.h, seeded
 * keys) when get_{ksk,bsk}_gpu first needs the key there. */
uint64_t concrete_b200_context_add_seeded_ksk(void *ctx, const uint64_t *bodies,
                                              uint64_t full_len_words, uint64_t seed_lo,
                                              uint64_t seed_hi);
uint64_t concrete_b200_context_add_seeded_bsk(void *ctx, const uint64_t *bodies,
                                              uint64_t full_len_words, uint64_t seed_lo,
                                              uint64_t seed_hi);
void concrete_b200_context_set_num_gpus(void *ctx, int n); /* same knob as env SDFG_NUM_GPUS */
void concrete_b200_context_set_device_base(void *ctx, int base); /* first device (LOCAL_RANK) */

/* wrappers.cpp:110-162 */
void memref_batched_keyswitch_lwe_cuda_u64(
    uint64_t *out_allocated, uint64_t *out_aligned, uint64_t out_offset, uint64_t out_size0,
    uint64_t out_size1, uint64_t out_stride0, uint64_t out_stride1, uint64_t *ct0_allocated,
    uint64_t *ct0_aligned, uint64_t ct0_offset, uint64_t ct0_size0, uint64_t ct0_size1,
    uint64_t ct0_stride0, uint64_t ct0_stride1, uint32_t level, uint32_t base_log,
    uint32_t input_lwe_dim, uint32_t output_lwe_dim, uint32_t ksk_index,
    cb_runtime_context *context);
/* wrappers.cpp:164-256 */
void memref_batched_bootstrap_lwe_cuda_u64(
    uint64_t *out_allocated, uint64_t *out_aligned, uint64_t out_offset, uint64_t out_size0,
    uint64_t out_size1, uint64_t out_stride0, uint64_t out_stride1, uint64_t *ct0_allocated,
    uint64_t *ct0_aligned, uint64_t ct0_offset, uint64_t ct0_size0, uint64_t ct0_size1,
    uint64_t ct0_stride0, uint64_t ct0_stride1, uint64_t *tlu_allocated, uint64_t *tlu_aligned,
    uint64_t tlu_offset, uint64_t tlu_size, uint64_t tlu_stride, uint32_t input_lwe_dim,
    uint32_t poly_size, uint32_t level, uint32_t base_log, uint32_t glwe_dim, uint32_t bsk_index,
    cb_runtime_context *context);
/* wrappers.cpp:258-363 */
void memref_batched_mapped_bootstrap_lwe_cuda_u64(
    uint64_t *out_allocated, uint64_t *out_aligned, uint64_t out_offset, uint64_t out_size0,
    uint64_t out_size1, uint64_t out_stride0, uint64_t out_stride1, uint64_t *ct0_allocated,
    uint64_t *ct0_aligned, uint64_t ct0_offset, uint64_t ct0_size0, uint64_t ct0_size1,
    uint64_t ct0_stride0, uint64_t ct0_stride1, uint64_t *tlu_allocated, uint64_t *tlu_aligned,
    uint64_t tlu_offset, uint64_t tlu_size0, uint64_t tlu_size1, uint64_t tlu_stride0,
    uint64_t tlu_stride1, uint32_t input_lwe_dim, uint32_t poly_size, uint32_t level,
    uint32_t base_log, uint32_t glwe_dim, uint32_t bsk_index, cb_runtime_context *context);
/* wrappers.cpp:70-106 */
void memref_keyswitch_lwe_cuda_u64(uint64_t *out_allocated, uint64_t *out_aligned,
                                   uint64_t out_offset, uint64_t out_size, uint64_t out_stride,
                                   uint64_t *ct0_allocated, uint64_t *ct0_aligned,
                                   uint64_t ct0_offset, uint64_t ct0_size, uint64_t ct0_stride,
                                   uint32_t level, uint32_t base_log, uint32_t input_lwe_dim,
                                   uint32_t output_lwe_dim, uint32_t ksk_index,
                                   cb_runtime_context *context);
void memref_bootstrap_lwe_cuda_u64(uint64_t *out_allocated, uint64_t *out_aligned,
                                   uint64_t out_offset, uint64_t out_size, uint64_t out_stride,
                                   uint64_t *ct0_allocated, uint64_t *ct0_aligned,
                                   uint64_t ct0_offset, uint64_t ct0_size, uint64_t ct0_stride,
                                   uint64_t *tlu_allocated, uint64_t *tlu_aligned,
                                   uint64_t tlu_offset, uint64_t tlu_size, uint64_t tlu_stride,
                                   uint32_t input_lwe_dim, uint32_t poly_size, uint32_t level,
                                   uint32_t base_log, uint32_t glwe_dim, uint32_t bsk_index,
                                   cb_runtime_context *context);
/* wrappers.cpp:745-760, 807-826, 828-849: CPU-named wrappers, same prototypes, GPU path here */
void memref_batched_keyswitch_lwe_u64(
    uint64_t *out_allocated, uint64_t *out_aligned, uint64_t out_offset, uint64_t out_size0,
    uint64_t out_size1, uint64_t out_stride0, uint64_t out_stride1, uint64_t *ct0_allocated,
    uint64_t *ct0_aligned, uint64_t ct0_offset, uint64_t ct0_size0, uint64_t ct0_size1,
    uint64_t ct0_stride0, uint64_t ct0_stride1, uint32_t level, uint32_t base_log,
    uint32_t input_lwe_dim, uint32_t output_lwe_dim, uint32_t ksk_index,
    cb_runtime_context *context);
void memref_batched_bootstrap_lwe_u64(
    uint64_t *out_allocated, uint64_t *out_aligned, uint64_t out_offset, uint64_t out_size0,
    uint64_t out_size1, uint64_t out_stride0, uint64_t out_stride1, uint64_t *ct0_allocated,
    uint64_t *ct0_aligned, uint64_t ct0_offset, uint64_t ct0_size0, uint64_t ct0_size1,
    uint64_t ct0_stride0, uint64_t ct0_stride1, uint64_t *tlu_allocated, uint64_t *tlu_aligned,
    uint64_t tlu_offset, uint64_t tlu_size, uint64_t tlu_stride, uint32_t input_lwe_dim,
    uint32_t poly_size, uint32_t level, uint32_t base_log, uint32_t glwe_dim, uint32_t bsk_index,
    cb_runtime_context *context);
void memref_batched_mapped_bootstrap_lwe_u64(
    uint64_t *out_allocated, uint64_t *out_aligned, uint64_t out_offset, uint64_t out_size0,
    uint64_t out_size1, uint64_t out_stride0, uint64_t out_stride1, uint64_t *ct0_allocated,
    uint64_t *ct0_aligned, uint64_t ct0_offset, uint64_t ct0_size0, uint64_t ct0_size1,
    uint64_t ct0_stride0, uint64_t ct0_stride1, uint64_t *tlu_allocated, uint64_t *tlu_aligned,
    uint64_t tlu_offset, uint64_t tlu_size0, uint64_t tlu_size1, uint64_t tlu_stride0,
    uint64_t tlu_stride1, uint32_t input_lwe_dim, uint32_t poly_size, uint32_t level,
    uint32_t base_log, uint32_t glwe_dim, uint32_t bsk_index, cb_runtime_context *context);
/* extension: KS -> PBS with intermediates kept on the device (GPUDFG.cpp:1085,1209 behaviour) */
void memref_batched_keyswitch_bootstrap_lwe_cuda_u64(
    uint64_t *out_aligned, uint64_t out_offset, uint64_t out_size0, uint64_t out_size1,
    uint64_t *ct0_aligned, uint64_t ct0_offset, uint64_t ct0_size0, uint64_t ct0_size1,
    uint64_t *tlu_aligned, uint64_t tlu_offset, uint64_t tlu_size, uint32_t ks_level,
    uint32_t ks_base_log, uint32_t big_lwe_dim, uint32_t small_lwe_dim, uint32_t ksk_index,
    uint32_t poly_size, uint32_t level, uint32_t base_log, uint32_t glwe_dim, uint32_t bsk_index,
    cb_runtime_context *context);
/* WoP-PBS over CRT blocks with HOST memrefs: the reference's CPU-only wrapper
 * (compilers/concrete-compiler/compiler/lib/Runtime/wrappers.cpp:855-997), same prototype, run on
 * the device (csrc/wop_pbs.cu); `pksk_index` selects a key registered with
 * concrete_b200_context_add_pksk (context.h fp_keyswitch_key_buffer). */
uint64_t concrete_b200_context_add_pksk(void *ctx, const uint64_t *data, uint64_t len_words);
void memref_wop_pbs_crt_buffer(
    uint64_t *out_allocated, uint64_t *out_aligned, uint64_t out_offset, uint64_t out_size_0,
    uint64_t out_size_1, uint64_t out_stride_0, uint64_t out_stride_1, uint64_t *in_allocated,
    uint64_t *in_aligned, uint64_t in_offset, uint64_t in_size_0, uint64_t in_size_1,
    uint64_t in_stride_0, uint64_t in_stride_1, uint64_t *lut_ct_allocated,
    uint64_t *lut_ct_aligned, uint64_t lut_ct_offset, uint64_t lut_ct_size0, uint64_t lut_ct_size1,
    uint64_t lut_ct_stride0, uint64_t lut_ct_stride1, uint64_t *crt_decomp_allocated,
    uint64_t *crt_decomp_aligned, uint64_t crt_decomp_offset, uint64_t crt_decomp_size,
    uint64_t crt_decomp_stride, uint32_t lwe_small_dim, uint32_t cbs_level_count,
    uint32_t cbs_base_log, uint32_t ksk_level_count, uint32_t ksk_base_log,
    uint32_t bsk_level_count, uint32_t bsk_base_log, uint32_t fpksk_level_count,
    uint32_t fpksk_base_log, uint32_t polynomial_size, uint32_t ksk_index, uint32_t bsk_index,
    uint32_t pksk_index, cb_runtime_context *context);

#ifdef __cplusplus
}
#endif
#endif
