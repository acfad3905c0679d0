"""The C-ABI library loads and exports every symbol include/*.h declares (no compute calls: this
runs without a GPU), and the ctypes table mirrors the headers."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    names = []
    for hdr in ("concrete_b200.h", "concrete_b200_runtime.h", "concrete_b200_sdfg.h"):
        text = open(os.path.join(ROOT, "include", hdr)).read()
        text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
        text = re.sub(r"//[^\n]*", "", text)
        for m in re.finditer(r"\b([A-Za-z_][A-Za-z0-9_]*)\s*\(", text):
            name = m.group(1)
            if name.startswith(("cuda_", "scratch_", "cleanup_", "memref_", "concrete_b200_", "stream_emulator_")):
                names.append(name)
    return sorted(set(names))


def test_headers_declare_the_reference_entry_points():
    names = _declared_symbols()
    for required in (
        "cuda_keyswitch_lwe_ciphertext_vector_64",
        "cuda_programmable_bootstrap_lwe_ciphertext_vector_64",
        "scratch_cuda_programmable_bootstrap_64",
        "cleanup_cuda_programmable_bootstrap",
        "cuda_convert_lwe_programmable_bootstrap_key_64",
        "cuda_programmable_bootstrap_amortized_lwe_ciphertext_vector_64",
        "cuda_bootstrap_low_latency_lwe_ciphertext_vector_64",
        "cuda_bootstrap_amortized_lwe_ciphertext_vector_64",
        "memref_batched_keyswitch_lwe_u64",
        "memref_batched_bootstrap_lwe_u64",
        "memref_batched_keyswitch_lwe_cuda_u64",
        "memref_batched_bootstrap_lwe_cuda_u64",
        "memref_batched_mapped_bootstrap_lwe_cuda_u64",
        "cuda_add_lwe_ciphertext_vector_64",
        # the SDFG code-generation interface (stream_emulator_api.h:29-104)
        "stream_emulator_init",
        "stream_emulator_run",
        "stream_emulator_delete",
        "stream_emulator_make_memref_add_lwe_ciphertexts_u64_process",
        "stream_emulator_make_memref_add_plaintext_lwe_ciphertext_u64_process",
        "stream_emulator_make_memref_mul_cleartext_lwe_ciphertext_u64_process",
        "stream_emulator_make_memref_negate_lwe_ciphertext_u64_process",
        "stream_emulator_make_memref_keyswitch_lwe_u64_process",
        "stream_emulator_make_memref_bootstrap_lwe_u64_process",
        "stream_emulator_make_memref_batched_add_lwe_ciphertexts_u64_process",
        "stream_emulator_make_memref_batched_add_plaintext_lwe_ciphertext_u64_process",
        "stream_emulator_make_memref_batched_add_plaintext_cst_lwe_ciphertext_u64_process",
        "stream_emulator_make_memref_batched_mul_cleartext_lwe_ciphertext_u64_process",
        "stream_emulator_make_memref_batched_mul_cleartext_cst_lwe_ciphertext_u64_process",
        "stream_emulator_make_memref_batched_negate_lwe_ciphertext_u64_process",
        "stream_emulator_make_memref_batched_keyswitch_lwe_u64_process",
        "stream_emulator_make_memref_batched_bootstrap_lwe_u64_process",
        "stream_emulator_make_memref_batched_mapped_bootstrap_lwe_u64_process",
        "stream_emulator_make_uint64_stream",
        "stream_emulator_put_uint64",
        "stream_emulator_get_uint64",
        "stream_emulator_make_memref_stream",
        "stream_emulator_put_memref",
        "stream_emulator_get_memref",
        "stream_emulator_make_memref_batch_stream",
        "stream_emulator_put_memref_batch",
        "stream_emulator_get_memref_batch",
    ):
        assert required in names, required


def test_library_exports_every_declared_symbol():
    from concrete_b200 import build as cb_build

    path = cb_build.build()
    lib = ctypes.CDLL(path, mode=ctypes.RTLD_LOCAL)
    missing = [n for n in _declared_symbols() if not hasattr(lib, n)]
    assert not missing, missing


def test_ctypes_table_covers_the_headers():
    from concrete_b200.lib import SIGNATURES, load_library

    declared = set(_declared_symbols())
    assert declared <= set(SIGNATURES), sorted(declared - set(SIGNATURES))
    lib = load_library()  # binds every signature; AttributeError if one is not exported
    assert lib.concrete_b200_version().startswith(b"concrete_b200")


def test_argument_counts_match_the_headers():
    """Arity of each ctypes prototype equals the number of parameters in the header."""
    from concrete_b200.lib import SIGNATURES

    for hdr in ("concrete_b200.h", "concrete_b200_runtime.h", "concrete_b200_sdfg.h"):
        text = open(os.path.join(ROOT, "include", hdr)).read()
        text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
        for m in re.finditer(r"\b([A-Za-z_][A-Za-z0-9_]*)\s*\(([^;{]*?)\)\s*;", text, flags=re.S):
            name, params = m.group(1), m.group(2).strip()
            if name not in SIGNATURES:
                continue
            n = 0 if params in ("", "void") else params.count(",") + 1
            assert n == len(SIGNATURES[name][1]), (name, n, len(SIGNATURES[name][1]))


def test_product_package_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "concrete_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in text and "from oracle" not in text, f
                assert "tfhe_oracle" not in text and "liboracle" not in text, f


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    import concrete_b200.lib as cl

    monkeypatch.setattr(cl, "_LIB", None)
    monkeypatch.setattr(cl, "_HERE", str(tmp_path))
    try:
        cl.load_library()
    except RuntimeError as e:
        assert "no CPU fallback" in str(e)
    else:
        raise AssertionError("load_library() must raise when the .so is missing")
