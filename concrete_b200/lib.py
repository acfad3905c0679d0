"""ctypes declaration of every symbol in include/concrete_b200.h (the C-ABI boundary)."""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

vp, u32, u64, i8pp = C.c_void_p, C.c_uint32, C.c_uint64, C.POINTER(C.POINTER(C.c_int8))
i8p = C.POINTER(C.c_int8)

# name -> (restype, argtypes); mirrors include/concrete_b200.h one to one
SIGNATURES = {
    "cuda_create_stream": (vp, [u32]),
    "cuda_destroy_stream": (None, [vp, u32]),
    "cuda_synchronize_stream": (None, [vp, u32]),
    "cuda_malloc": (vp, [u64, u32]),
    "cuda_malloc_async": (vp, [u64, vp, u32]),
    "cuda_check_valid_malloc": (None, [u64, u32]),
    "cuda_memcpy_async_to_gpu": (None, [vp, vp, u64, vp, u32]),
    "cuda_memcpy_async_gpu_to_gpu": (None, [vp, vp, u64, vp, u32]),
    "cuda_memcpy_gpu_to_gpu": (None, [vp, vp, u64, u32]),
    "cuda_memcpy_async_to_cpu": (None, [vp, vp, u64, vp, u32]),
    "cuda_memset_async": (None, [vp, u64, u64, vp, u32]),
    "cuda_get_number_of_gpus": (C.c_int, []),
    "cuda_synchronize_device": (None, [u32]),
    "cuda_drop": (None, [vp, u32]),
    "cuda_drop_async": (None, [vp, vp, u32]),
    "cuda_get_max_shared_memory": (C.c_int, [u32]),
    "cuda_keyswitch_lwe_ciphertext_vector_64": (None, [vp, u32, vp, vp, vp, vp, vp, u32, u32, u32, u32, u32]),
    "concrete_b200_keyswitch_lwe_ciphertext_vector_64_no_alloc": (
        C.c_int, [vp, u32, vp, vp, vp, vp, vp, u32, u32, u32, u32, u32, C.c_char_p, u64]),
    "cuda_convert_lwe_programmable_bootstrap_key_64": (None, [vp, u32, vp, vp, u32, u32, u32, u32]),
    "scratch_cuda_programmable_bootstrap_64": (None, [vp, u32, i8pp, u32, u32, u32, u32, C.c_bool]),
    "cuda_programmable_bootstrap_lwe_ciphertext_vector_64": (
        None, [vp, u32, vp, vp, vp, vp, vp, vp, vp, i8p, u32, u32, u32, u32, u32, u32, u32, u32]),
    "cleanup_cuda_programmable_bootstrap": (None, [vp, u32, i8pp]),
    "scratch_cuda_programmable_bootstrap_amortized_64": (None, [vp, u32, i8pp, u32, u32, u32, C.c_bool]),
    "cuda_programmable_bootstrap_amortized_lwe_ciphertext_vector_64": (
        None, [vp, u32, vp, vp, vp, vp, vp, vp, vp, i8p, u32, u32, u32, u32, u32, u32]),
    "cleanup_cuda_programmable_bootstrap_amortized": (None, [vp, u32, i8pp]),
    "cuda_fourier_polynomial_mul": (None, [vp, u32, vp, vp, vp, u32, u32]),
    "cuda_negate_lwe_ciphertext_vector_64": (None, [vp, u32, vp, vp, u32, u32]),
    "cuda_add_lwe_ciphertext_vector_64": (None, [vp, u32, vp, vp, vp, u32, u32]),
    "cuda_add_lwe_ciphertext_vector_plaintext_vector_64": (None, [vp, u32, vp, vp, vp, u32, u32]),
    "cuda_mult_lwe_ciphertext_vector_cleartext_vector_64": (None, [vp, u32, vp, vp, vp, u32, u32]),
    "cuda_convert_lwe_bootstrap_key_64": (None, [vp, vp, vp, u32, u32, u32, u32, u32]),
    "scratch_cuda_bootstrap_low_latency_64": (None, [vp, u32, i8pp, u32, u32, u32, u32, u32, C.c_bool]),
    "cuda_bootstrap_low_latency_lwe_ciphertext_vector_64": (
        None, [vp, u32, vp, vp, vp, vp, vp, i8p, u32, u32, u32, u32, u32, u32, u32, u32, u32]),
    "cleanup_cuda_bootstrap_low_latency": (None, [vp, u32, i8pp]),
    "scratch_cuda_bootstrap_amortized_64": (None, [vp, u32, i8pp, u32, u32, u32, u32, C.c_bool]),
    "cuda_bootstrap_amortized_lwe_ciphertext_vector_64": (
        None, [vp, u32, vp, vp, vp, vp, vp, i8p, u32, u32, u32, u32, u32, u32, u32, u32, u32]),
    "cleanup_cuda_bootstrap_amortized": (None, [vp, u32, i8pp]),
    "cuda_keyswitch_lwe_ciphertext_vector_64_legacy": (None, [vp, u32, vp, vp, vp, u32, u32, u32, u32, u32]),
    "cuda_keyswitch_bootstrap_lwe_ciphertext_vector_64": (
        None, [vp, u32, vp, vp, vp, vp, vp, vp, i8p, vp, u32, u32, u32, u32, u32, u32, u32, u32, u32]),
    "cuda_decompress_seeded_lwe_bootstrap_key_64": (None, [vp, u32, vp, vp, u32, u32, u32, u32, u64, u64]),
    "cuda_decompress_seeded_lwe_keyswitch_key_64": (None, [vp, u32, vp, vp, u32, u32, u32, u64, u64]),
    "cuda_convert_seeded_lwe_programmable_bootstrap_key_64": (
        None, [vp, u32, vp, vp, u32, u32, u32, u32, u64, u64]),
    "cuda_extract_bits_lwe_ciphertext_vector_64": (
        None, [vp, u32, vp, vp, vp, u32, vp, vp, u32, u32, u32, u32, u32, u32, u32]),
    "cuda_circuit_bootstrap_vertical_packing_lwe_ciphertext_vector_64": (
        None, [vp, u32, vp, vp, vp, vp, vp, u32, u32, u64, u32, u32, u32, u32, u32, u32, u32, u32, u32]),
    "cuda_wop_pbs_crt_lwe_ciphertext_vector_64": (
        None, [vp, u32, vp, vp, vp, vp, u32, vp, vp, vp, u32, u32, u32, u32, u32, u32, u32, u32, u32,
               u32, u32]),
    "concrete_b200_version": (C.c_char_p, []),
    "concrete_b200_kernel_launches": (u64, []),
    "concrete_b200_set_pbs_plan": (None, [C.c_int]),
    "concrete_b200_set_ks_plan": (None, [C.c_int]),
    "concrete_b200_set_v4_alt": (None, [C.c_int]),
}

# runtime wrapper layer, include/concrete_b200_runtime.h (MLIR memref ABI; host pointers)
_m2 = [vp, vp, u64, u64, u64, u64, u64]  # 2-D memref
_m1 = [vp, vp, u64, u64, u64]            # 1-D memref
SIGNATURES.update({
    "concrete_b200_context_create": (vp, []),
    "concrete_b200_context_destroy": (None, [vp]),
    "concrete_b200_context_add_ksk": (u64, [vp, vp, u64]),
    "concrete_b200_context_add_bsk": (u64, [vp, vp, u64]),
    "concrete_b200_context_add_pksk": (u64, [vp, vp, u64]),
    "memref_wop_pbs_crt_buffer": (None, [vp, vp, u64, u64, u64, u64, u64, vp, vp, u64, u64, u64, u64, u64,
                                         vp, vp, u64, u64, u64, u64, u64, vp, vp, u64, u64, u64] + [u32] * 13 + [vp]),
    "concrete_b200_context_add_seeded_ksk": (u64, [vp, vp, u64, u64, u64]),
    "concrete_b200_context_add_seeded_bsk": (u64, [vp, vp, u64, u64, u64]),
    "concrete_b200_context_set_num_gpus": (None, [vp, C.c_int]),
    "concrete_b200_context_set_device_base": (None, [vp, C.c_int]),
    "memref_batched_keyswitch_lwe_cuda_u64": (None, _m2 + _m2 + [u32, u32, u32, u32, u32, vp]),
    "memref_batched_bootstrap_lwe_cuda_u64": (None, _m2 + _m2 + _m1 + [u32, u32, u32, u32, u32, u32, vp]),
    "memref_batched_mapped_bootstrap_lwe_cuda_u64": (None, _m2 + _m2 + _m2 + [u32, u32, u32, u32, u32, u32, vp]),
    "memref_keyswitch_lwe_cuda_u64": (None, _m1 + _m1 + [u32, u32, u32, u32, u32, vp]),
    "memref_bootstrap_lwe_cuda_u64": (None, _m1 + _m1 + _m1 + [u32, u32, u32, u32, u32, u32, vp]),
    "memref_batched_keyswitch_lwe_u64": (None, _m2 + _m2 + [u32, u32, u32, u32, u32, vp]),
    "memref_batched_bootstrap_lwe_u64": (None, _m2 + _m2 + _m1 + [u32, u32, u32, u32, u32, u32, vp]),
    "memref_batched_mapped_bootstrap_lwe_u64": (None, _m2 + _m2 + _m2 + [u32, u32, u32, u32, u32, u32, vp]),
    "memref_batched_keyswitch_bootstrap_lwe_cuda_u64": (
        None, [vp, u64, u64, u64, vp, u64, u64, u64, vp, u64, u64] + [u32] * 10 + [vp]),
    "concrete_b200_fp64_peak": (None, [vp, u32, vp, u32, u32]),
})

# SDFG ("stream emulator") code-generation interface, include/concrete_b200_runtime.h
# (stream_emulator_api.h:29-104 of the reference)
_p4 = [vp, vp, vp, vp]
_ks = [vp, vp, vp, u32, u32, u32, u32, u32, u32, vp]
_bs = [vp, vp, vp, vp, u32, u32, u32, u32, u32, u32, u32, vp]
SIGNATURES.update({
    "stream_emulator_init": (vp, []),
    "stream_emulator_run": (None, [vp]),
    "stream_emulator_delete": (None, [vp]),
    "stream_emulator_make_memref_add_lwe_ciphertexts_u64_process": (None, _p4),
    "stream_emulator_make_memref_add_plaintext_lwe_ciphertext_u64_process": (None, _p4),
    "stream_emulator_make_memref_mul_cleartext_lwe_ciphertext_u64_process": (None, _p4),
    "stream_emulator_make_memref_negate_lwe_ciphertext_u64_process": (None, [vp, vp, vp]),
    "stream_emulator_make_memref_keyswitch_lwe_u64_process": (None, _ks),
    "stream_emulator_make_memref_bootstrap_lwe_u64_process": (None, _bs),
    "stream_emulator_make_memref_batched_add_lwe_ciphertexts_u64_process": (None, _p4),
    "stream_emulator_make_memref_batched_add_plaintext_lwe_ciphertext_u64_process": (None, _p4),
    "stream_emulator_make_memref_batched_add_plaintext_cst_lwe_ciphertext_u64_process": (None, _p4),
    "stream_emulator_make_memref_batched_mul_cleartext_lwe_ciphertext_u64_process": (None, _p4),
    "stream_emulator_make_memref_batched_mul_cleartext_cst_lwe_ciphertext_u64_process": (None, _p4),
    "stream_emulator_make_memref_batched_negate_lwe_ciphertext_u64_process": (None, [vp, vp, vp]),
    "stream_emulator_make_memref_batched_keyswitch_lwe_u64_process": (None, _ks),
    "stream_emulator_make_memref_batched_bootstrap_lwe_u64_process": (None, _bs),
    "stream_emulator_make_memref_batched_mapped_bootstrap_lwe_u64_process": (None, _bs),
    "stream_emulator_make_uint64_stream": (vp, [C.c_char_p, C.c_int]),
    "stream_emulator_put_uint64": (None, [vp, u64]),
    "stream_emulator_get_uint64": (u64, [vp]),
    "stream_emulator_make_memref_stream": (vp, [C.c_char_p, C.c_int]),
    "stream_emulator_put_memref": (None, [vp, vp, vp, u64, u64, u64, u64]),
    "stream_emulator_get_memref": (None, [vp, vp, vp, u64, u64, u64]),
    "stream_emulator_make_memref_batch_stream": (vp, [C.c_char_p, C.c_int]),
    "stream_emulator_put_memref_batch": (None, [vp, vp, vp, u64, u64, u64, u64, u64, u64]),
    "stream_emulator_get_memref_batch": (None, [vp, vp, vp, u64, u64, u64, u64, u64]),
    "concrete_b200_sdfg_set_num_gpus": (None, [vp, C.c_int]),
    "concrete_b200_sdfg_delete_stream": (None, [vp]),
})


def library_path() -> str:
    return os.path.join(_HERE, "libconcrete_b200.so")


def load_library():
    """Load libconcrete_b200.so.  No fallback: a missing library is an error."""
    global _LIB
    if _LIB is not None:
        return _LIB
    path = library_path()
    if not os.path.exists(path):
        raise RuntimeError(
            f"{path} is missing: build it with `python -m concrete_b200.build` "
            "(nvcc, sm_100a).  concrete_b200 has no CPU fallback."
        )
    lib = C.CDLL(path, mode=C.RTLD_LOCAL)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    _LIB = lib
    return lib
