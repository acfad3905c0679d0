/* concrete_b200_sdfg.h — the SDFG ("stream emulator") code-generation interface on the B200 path.
 *
 * These are the symbols Concrete's generated code calls when a circuit is compiled with the SDFG
 * lowering (reference interface: compilers/concrete-compiler/compiler/include/concretelang/
 * Runtime/stream_emulator_api.h:29-104; reference implementation: lib/Runtime/GPUDFG.cpp:1467-1799).
 * Same names, argument order and meaning, so libconcrete_b200.so resolves them in place of
 * libConcretelangRuntime's GPUDFG.cpp.  Implementation: concrete_b200/csrc/sdfg.cpp.
 *
 *   dfg      = stream_emulator_init()
 *   stream   = stream_emulator_make_memref[_batch]_stream(name, type)
 *              stream_emulator_make_*_process(dfg, in streams..., out stream, parameters..., ctx)
 *              stream_emulator_put_memref[_batch](stream, host memref, ownership)
 *              stream_emulator_get_memref[_batch](stream, host memref)      <- evaluation happens here
 *              stream_emulator_delete(dfg)
 *
 * Differences from the reference scheduler, all invisible through this interface:
 *  - a GET evaluates the stale sub-graph feeding that stream; intermediates stay on the device
 *    (per-GPU shards) and are only downloaded if they are themselves the target of a GET;
 *  - each device runs its contiguous shard as a chunk pipeline on three streams
 *    (upload / compute / download overlap) with per-device persistent scratch;
 *  - the number of devices is SDFG_NUM_GPUS as in GPUDFG.cpp:1749-1755, or
 *    concrete_b200_sdfg_set_num_gpus().
 *  `context` is the RuntimeContext* the generated code passes through (runtime_context.h). */
#ifndef CONCRETE_B200_SDFG_H
#define CONCRETE_B200_SDFG_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* stream_emulator_api.h:16-22 */
typedef enum cb_stream_type {
  CB_TS_STREAM_TYPE_X86_TO_TOPO_LSAP = 0,
  CB_TS_STREAM_TYPE_TOPO_TO_TOPO_LSAP = 1,
  CB_TS_STREAM_TYPE_TOPO_TO_X86_LSAP = 2,
  CB_TS_STREAM_TYPE_TOPO_TO_BOTH_LSAP = 3,
  CB_TS_STREAM_TYPE_X86_TO_X86_LSAP = 4
} cb_stream_type;

/* graph lifetime — stream_emulator_api.h:30-32 */
void *stream_emulator_init(void);
void stream_emulator_run(void *dfg);
void stream_emulator_delete(void *dfg);

/* one-ciphertext processes — stream_emulator_api.h:34-53 */
void stream_emulator_make_memref_add_lwe_ciphertexts_u64_process(void *dfg, void *sin1, void *sin2, void *sout);
void stream_emulator_make_memref_add_plaintext_lwe_ciphertext_u64_process(void *dfg, void *sin1, void *sin2, void *sout);
void stream_emulator_make_memref_mul_cleartext_lwe_ciphertext_u64_process(void *dfg, void *sin1, void *sin2, void *sout);
void stream_emulator_make_memref_negate_lwe_ciphertext_u64_process(void *dfg, void *sin1, void *sout);
void stream_emulator_make_memref_keyswitch_lwe_u64_process(
    void *dfg, void *sin1, void *sout, uint32_t level, uint32_t base_log, uint32_t input_lwe_dim,
    uint32_t output_lwe_dim, uint32_t output_size, uint32_t ksk_index, void *context);
void stream_emulator_make_memref_bootstrap_lwe_u64_process(
    void *dfg, void *sin1, void *sin2, void *sout, uint32_t input_lwe_dim, uint32_t poly_size,
    uint32_t level, uint32_t base_log, uint32_t glwe_dim, uint32_t output_size, uint32_t bsk_index,
    void *context);

/* batched processes — stream_emulator_api.h:55-81.  `_cst` = second operand is one scalar;
 * otherwise a per-ciphertext vector presented as a {1, num} batch (GPUDFG.cpp:1306-1308). */
void stream_emulator_make_memref_batched_add_lwe_ciphertexts_u64_process(void *dfg, void *sin1, void *sin2, void *sout);
void stream_emulator_make_memref_batched_add_plaintext_lwe_ciphertext_u64_process(void *dfg, void *sin1, void *sin2, void *sout);
void stream_emulator_make_memref_batched_add_plaintext_cst_lwe_ciphertext_u64_process(void *dfg, void *sin1, void *sin2, void *sout);
void stream_emulator_make_memref_batched_mul_cleartext_lwe_ciphertext_u64_process(void *dfg, void *sin1, void *sin2, void *sout);
void stream_emulator_make_memref_batched_mul_cleartext_cst_lwe_ciphertext_u64_process(void *dfg, void *sin1, void *sin2, void *sout);
void stream_emulator_make_memref_batched_negate_lwe_ciphertext_u64_process(void *dfg, void *sin1, void *sout);
void stream_emulator_make_memref_batched_keyswitch_lwe_u64_process(
    void *dfg, void *sin1, void *sout, uint32_t level, uint32_t base_log, uint32_t input_lwe_dim,
    uint32_t output_lwe_dim, uint32_t output_size, uint32_t ksk_index, void *context);
void stream_emulator_make_memref_batched_bootstrap_lwe_u64_process(
    void *dfg, void *sin1, void *sin2, void *sout, uint32_t input_lwe_dim, uint32_t poly_size,
    uint32_t level, uint32_t base_log, uint32_t glwe_dim, uint32_t output_size, uint32_t bsk_index,
    void *context);
void stream_emulator_make_memref_batched_mapped_bootstrap_lwe_u64_process(
    void *dfg, void *sin1, void *sin2, void *sout, uint32_t input_lwe_dim, uint32_t poly_size,
    uint32_t level, uint32_t base_log, uint32_t glwe_dim, uint32_t output_size, uint32_t bsk_index,
    void *context);

/* streams — stream_emulator_api.h:83-103.  data_ownership != 0: the stream takes the host
 * buffer (freed with free() when replaced or deleted), as GPUDFG.cpp:1617-1650. */
void *stream_emulator_make_uint64_stream(const char *name, cb_stream_type stype);
void stream_emulator_put_uint64(void *stream, uint64_t e);
uint64_t stream_emulator_get_uint64(void *stream);
void *stream_emulator_make_memref_stream(const char *name, cb_stream_type stype);
void stream_emulator_put_memref(void *stream, uint64_t *allocated, uint64_t *aligned,
                                uint64_t offset, uint64_t size, uint64_t stride,
                                uint64_t data_ownership);
void stream_emulator_get_memref(void *stream, uint64_t *out_allocated, uint64_t *out_aligned,
                                uint64_t out_offset, uint64_t out_size, uint64_t out_stride);
void *stream_emulator_make_memref_batch_stream(const char *name, cb_stream_type stype);
void stream_emulator_put_memref_batch(void *stream, uint64_t *allocated, uint64_t *aligned,
                                      uint64_t offset, uint64_t size0, uint64_t size1,
                                      uint64_t stride0, uint64_t stride1, uint64_t data_ownership);
void stream_emulator_get_memref_batch(void *stream, uint64_t *out_allocated, uint64_t *out_aligned,
                                      uint64_t out_offset, uint64_t out_size0, uint64_t out_size1,
                                      uint64_t out_stride0, uint64_t out_stride1);

/* extensions (no reference counterpart) */
void concrete_b200_sdfg_set_num_gpus(void *dfg, int num_gpus); /* instead of SDFG_NUM_GPUS */
void concrete_b200_sdfg_delete_stream(void *stream); /* streams not attached to any process */

#ifdef __cplusplus
}
#endif
#endif
