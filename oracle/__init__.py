"""ctypes binding of the CPU oracle (oracle/tfhe_oracle.cpp).

TEST INFRASTRUCTURE ONLY: importable from tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / ``--impl reference`` legs.  The product package
``concrete_b200`` never imports this module.

Every function mirrors one C entry point; see the C++ header comment for the
reference file:line each one restates.  Arrays are numpy ``uint64`` (ciphertexts,
keys) or ``complex128`` (Fourier keys), C-contiguous.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None
_LIB_NATIVE = None

u64p = C.POINTER(C.c_uint64)


def _native_name() -> str:
    """-march=native objects must not travel between machines: key the file on the CPU flags."""
    import hashlib

    flags = ""
    try:
        with open("/proc/cpuinfo") as f:
            for ln in f:
                if ln.startswith("flags"):
                    flags = ln
                    break
    except OSError:
        pass
    return "liboracle_native_%s.so" % hashlib.sha1(flags.encode()).hexdigest()[:10]


def build(native: bool = False) -> str:
    """Compile the oracle with the committed Makefile; returns the .so path."""
    target = _native_name() if native else "liboracle.so"
    args = ["make", "-C", _HERE, target]
    if native:
        args = ["make", "-C", _HERE, "native", f"NATIVE_SO={target}"]
    subprocess.run(args, check=True, capture_output=True)
    return os.path.join(_HERE, target)


def lib(native: bool = False):
    global _LIB, _LIB_NATIVE
    if _LIB is not None and not native:
        return _LIB
    if _LIB_NATIVE is not None and native:
        return _LIB_NATIVE
    path = os.path.join(_HERE, _native_name() if native else "liboracle.so")
    srcs = [os.path.join(_HERE, f) for f in ("tfhe_oracle.cpp", "csprng_oracle.cpp", "wop_oracle.inl")]
    if not os.path.exists(path) or any(
        os.path.exists(src) and os.path.getmtime(src) > os.path.getmtime(path) for src in srcs
    ):
        build(native)
    l = C.CDLL(path)
    l.orc_closest_representable_u8.restype = C.c_uint8
    l.orc_closest_representable_u8.argtypes = [C.c_uint8, C.c_uint, C.c_uint]
    l.orc_closest_representable_u32.restype = C.c_uint32
    l.orc_closest_representable_u32.argtypes = [C.c_uint32, C.c_uint, C.c_uint]
    l.orc_closest_representable_u64.restype = C.c_uint64
    l.orc_closest_representable_u64.argtypes = [C.c_uint64, C.c_uint, C.c_uint]
    l.orc_decompose_u32.argtypes = [C.c_uint32, C.c_uint, C.c_uint, C.POINTER(C.c_uint32)]
    l.orc_decompose_u64.argtypes = [C.c_uint64, C.c_uint, C.c_uint, u64p]
    l.orc_modulus_switch.restype = C.c_uint64
    l.orc_modulus_switch.argtypes = [C.c_uint64, C.c_uint]
    l.orc_from_torus.restype = C.c_uint64
    l.orc_from_torus.argtypes = [C.c_double]
    l.orc_num_threads.restype = C.c_int
    l.orc_set_num_threads.argtypes = [C.c_int]
    if native:
        _LIB_NATIVE = l
    else:
        _LIB = l
    return l


_FAST = None


def fast_lib():
    """The vectorised timed baseline (oracle/tfhe_fast.cpp), -march=native, built where it runs."""
    global _FAST
    if _FAST is not None:
        return _FAST
    name = _native_name().replace("liboracle_native", "libfast_native")
    path = os.path.join(_HERE, name)
    src = os.path.join(_HERE, "tfhe_fast.cpp")
    if not os.path.exists(path) or os.path.getmtime(src) > os.path.getmtime(path):
        subprocess.run(["make", "-C", _HERE, "fast", f"FAST_SO={name}"], check=True,
                       capture_output=True)
    l = C.CDLL(path)
    l.fast_num_threads.restype = C.c_int
    l.fast_set_num_threads.argtypes = [C.c_int]
    l.fast_simd_bits.restype = C.c_int
    _FAST = l
    return l


def fast_keyswitch(cts, ksk, n_in, n_out, base_log, level) -> np.ndarray:
    num = cts.shape[0]
    out = np.empty((num, n_out + 1), dtype=np.uint64)
    fast_lib().fast_keyswitch_batch(_p(out), _p(cts), _p(ksk), n_in, n_out, base_log, level,
                                    C.c_size_t(num))
    return out


def fast_bootstrap(cts, luts, lut_idx, fbsk, n, k, N, base_log, level) -> np.ndarray:
    num = cts.shape[0]
    luts = np.ascontiguousarray(luts.reshape(-1, (k + 1) * N))
    out = np.empty((num, k * N + 1), dtype=np.uint64)
    idx = None if lut_idx is None else _p(np.ascontiguousarray(lut_idx, dtype=np.uint64))
    fast_lib().fast_bootstrap_batch(_p(out), _p(cts), _p(luts), idx, _p(fbsk, np.complex128), n, k,
                                    C.c_size_t(N), base_log, level, C.c_size_t(num))
    return out


def _p(a: np.ndarray, dtype=np.uint64):
    assert a.dtype == dtype and a.flags["C_CONTIGUOUS"], (a.dtype, a.flags)
    return a.ctypes.data_as(C.c_void_p)


def num_threads(native=False) -> int:
    return lib(native).orc_num_threads()


# ---------------------------------------------------------------- scalar pieces
def closest_representable(x: int, level: int, base_log: int, bits: int = 64) -> int:
    f = getattr(lib(), f"orc_closest_representable_u{bits}")
    return int(f(x, level, base_log))


def decompose(x: int, level: int, base_log: int, bits: int = 64):
    """Signed digits, index 0 = level `level` (least significant) ... last = level 1."""
    if bits == 64:
        out = (C.c_uint64 * level)()
        lib().orc_decompose_u64(x, level, base_log, out)
        return [v - (1 << 64) if v >> 63 else v for v in out]
    out = (C.c_uint32 * level)()
    lib().orc_decompose_u32(x, level, base_log, out)
    return [v - (1 << 32) if v >> 31 else v for v in out]


def modulus_switch(x: int, log_modulus: int) -> int:
    return int(lib().orc_modulus_switch(x, log_modulus))


def from_torus(x: float) -> int:
    return int(lib().orc_from_torus(x))


def monomial_div(poly: np.ndarray, degree: int) -> np.ndarray:
    out = np.empty_like(poly)
    if poly.dtype == np.uint8:
        lib().orc_monomial_div_u8(_p(out, np.uint8), _p(poly, np.uint8), C.c_size_t(len(poly)), C.c_size_t(degree))
    else:
        lib().orc_monomial_div_u64(_p(out), _p(poly), C.c_size_t(len(poly)), C.c_size_t(degree))
    return out


def monomial_mul(poly: np.ndarray, degree: int) -> np.ndarray:
    out = np.empty_like(poly)
    if poly.dtype == np.uint8:
        lib().orc_monomial_mul_u8(_p(out, np.uint8), _p(poly, np.uint8), C.c_size_t(len(poly)), C.c_size_t(degree))
    else:
        lib().orc_monomial_mul_u64(_p(out), _p(poly), C.c_size_t(len(poly)), C.c_size_t(degree))
    return out


def monomial_mul_and_subtract(poly: np.ndarray, degree: int) -> np.ndarray:
    out = np.empty_like(poly)
    lib().orc_monomial_mul_and_subtract_u64(_p(out), _p(poly), C.c_size_t(len(poly)), C.c_size_t(degree))
    return out


def sample_extract(glwe: np.ndarray, k: int, N: int, nth: int = 0) -> np.ndarray:
    out = np.empty(k * N + 1, dtype=np.uint64)
    lib().orc_sample_extract_u64(_p(out), _p(glwe), C.c_size_t(k), C.c_size_t(N), C.c_size_t(nth))
    return out


def negacyclic_mul_schoolbook(a, b):
    out = np.empty_like(a)
    lib().orc_negacyclic_mul_schoolbook(_p(out), _p(a), _p(b), C.c_size_t(len(a)))
    return out


def negacyclic_mul_binary(a, s):
    out = np.empty_like(a)
    lib().orc_negacyclic_mul_binary(_p(out), _p(a), _p(s), C.c_size_t(len(a)))
    return out


# ---------------------------------------------------------------- FFT pieces
def fft_forward_as_torus(poly: np.ndarray) -> np.ndarray:
    N = len(poly)
    out = np.empty(N // 2, dtype=np.complex128)
    lib().orc_fft_forward_as_torus(_p(out, np.complex128), _p(poly), C.c_size_t(N))
    return out


def fft_forward_as_integer(poly: np.ndarray) -> np.ndarray:
    N = len(poly)
    out = np.empty(N // 2, dtype=np.complex128)
    lib().orc_fft_forward_as_integer(_p(out, np.complex128), _p(poly), C.c_size_t(N))
    return out


def fft_add_backward_as_torus(poly: np.ndarray, fourier: np.ndarray) -> None:
    lib().orc_fft_add_backward_as_torus(_p(poly), _p(fourier, np.complex128), C.c_size_t(len(poly)))


# ---------------------------------------------------------------- hot path
def keyswitch(cts: np.ndarray, ksk: np.ndarray, n_in, n_out, base_log, level, native=False) -> np.ndarray:
    num = cts.shape[0]
    assert cts.shape[1] == n_in + 1 and ksk.size == n_in * level * (n_out + 1)
    out = np.empty((num, n_out + 1), dtype=np.uint64)
    lib(native).orc_keyswitch_batch(_p(out), _p(cts), _p(ksk), n_in, n_out, base_log, level, C.c_size_t(num))
    return out


def bsk_to_fourier(bsk: np.ndarray, n, k, level, N, native=False) -> np.ndarray:
    assert bsk.size == n * level * (k + 1) ** 2 * N
    out = np.empty(bsk.size // 2, dtype=np.complex128)
    lib(native).orc_bsk_to_fourier(_p(out, np.complex128), _p(bsk), n, k, level, C.c_size_t(N))
    return out


def bootstrap(cts, luts, lut_idx, fbsk, n, k, N, base_log, level, nth=0, native=False) -> np.ndarray:
    num = cts.shape[0]
    assert cts.shape[1] == n + 1
    luts = np.ascontiguousarray(luts.reshape(-1, (k + 1) * N))
    out = np.empty((num, k * N + 1), dtype=np.uint64)
    idx = None if lut_idx is None else _p(np.ascontiguousarray(lut_idx, dtype=np.uint64))
    lib(native).orc_bootstrap_batch(
        _p(out), _p(cts), _p(luts), idx, _p(fbsk, np.complex128), n, k, C.c_size_t(N),
        base_log, level, C.c_size_t(num), C.c_size_t(nth))
    return out


def blind_rotate(cts, luts, lut_idx, fbsk, n, k, N, base_log, level) -> np.ndarray:
    num = cts.shape[0]
    luts = np.ascontiguousarray(luts.reshape(-1, (k + 1) * N))
    out = np.empty((num, (k + 1) * N), dtype=np.uint64)
    idx = None if lut_idx is None else _p(np.ascontiguousarray(lut_idx, dtype=np.uint64))
    lib().orc_blind_rotate_batch(
        _p(out), _p(cts), _p(luts), idx, _p(fbsk, np.complex128), n, k, C.c_size_t(N),
        base_log, level, C.c_size_t(num))
    return out


# ---------------------------------------------------------------- harness
def gen_binary_key(length: int, seed: int) -> np.ndarray:
    key = np.empty(length, dtype=np.uint64)
    lib().orc_gen_binary_key(_p(key), C.c_size_t(length), C.c_uint64(seed))
    return key


def lwe_encrypt(key: np.ndarray, plaintexts: np.ndarray, sigma: float, seed: int) -> np.ndarray:
    n, num = len(key), len(plaintexts)
    ct = np.empty((num, n + 1), dtype=np.uint64)
    lib().orc_lwe_encrypt_batch(_p(ct), _p(key), _p(np.ascontiguousarray(plaintexts, dtype=np.uint64)),
                                n, C.c_size_t(num), C.c_double(sigma), C.c_uint64(seed))
    return ct


def lwe_decrypt(key: np.ndarray, cts: np.ndarray) -> np.ndarray:
    n, num = len(key), cts.shape[0]
    assert cts.shape[1] == n + 1
    pt = np.empty(num, dtype=np.uint64)
    lib().orc_lwe_decrypt_batch(_p(pt), _p(np.ascontiguousarray(cts)), _p(key), n, C.c_size_t(num))
    return pt


def gen_ksk(key_in, key_out, base_log, level, sigma, seed) -> np.ndarray:
    n_in, n_out = len(key_in), len(key_out)
    ksk = np.empty(n_in * level * (n_out + 1), dtype=np.uint64)
    lib().orc_gen_ksk(_p(ksk), _p(key_in), _p(key_out), n_in, n_out, base_log, level,
                      C.c_double(sigma), C.c_uint64(seed))
    return ksk


def gen_bsk(lwe_key, glwe_key, k, N, base_log, level, sigma, seed) -> np.ndarray:
    n = len(lwe_key)
    assert len(glwe_key) == k * N
    bsk = np.empty(n * level * (k + 1) ** 2 * N, dtype=np.uint64)
    lib().orc_gen_bsk(_p(bsk), _p(lwe_key), _p(glwe_key), n, k, C.c_size_t(N), base_log, level,
                      C.c_double(sigma), C.c_uint64(seed))
    return bsk


def gen_ksk_masks(key_in, key_out, base_log, level, sigma, seed, masks) -> np.ndarray:
    """gen_ksk with the mask elements given ([n_in][level][n_out]): seeded-key tests."""
    n_in, n_out = len(key_in), len(key_out)
    assert masks.size == n_in * level * n_out
    ksk = np.empty(n_in * level * (n_out + 1), dtype=np.uint64)
    lib().orc_gen_ksk_masks(_p(ksk), _p(key_in), _p(key_out), n_in, n_out, base_log, level,
                            C.c_double(sigma), C.c_uint64(seed), _p(np.ascontiguousarray(masks)))
    return ksk


def gen_bsk_masks(lwe_key, glwe_key, k, N, base_log, level, sigma, seed, masks) -> np.ndarray:
    """gen_bsk with the mask polynomials given ([n][level][k+1][k][N]): seeded-key tests."""
    n = len(lwe_key)
    assert masks.size == n * level * (k + 1) * k * N
    bsk = np.empty(n * level * (k + 1) ** 2 * N, dtype=np.uint64)
    lib().orc_gen_bsk_masks(_p(bsk), _p(lwe_key), _p(glwe_key), n, k, C.c_size_t(N), base_log, level,
                            C.c_double(sigma), C.c_uint64(seed), _p(np.ascontiguousarray(masks)))
    return bsk


# ---------------------------------------------------------------- seeded keys (csprng_oracle.cpp)
def _seed_halves(seed: int):
    return C.c_uint64(seed & (2 ** 64 - 1)), C.c_uint64(seed >> 64)


def aes128_encrypt(key: bytes, block: bytes) -> bytes:
    out = (C.c_uint8 * 16)()
    lib().orc_aes128_encrypt(key, block, out)
    return bytes(out)


def csprng_bytes(seed: int, first_byte: int, count: int) -> np.ndarray:
    """Bytes [first_byte, first_byte + count) a fresh generator of `seed` (u128) outputs."""
    out = np.empty(count, dtype=np.uint8)
    lo, hi = _seed_halves(seed)
    lib().orc_csprng_bytes(_p(out, np.uint8), C.c_uint64(first_byte), C.c_uint64(count), lo, hi)
    return out


def csprng_u64(seed: int, first_word: int, count: int) -> np.ndarray:
    out = np.empty(count, dtype=np.uint64)
    lo, hi = _seed_halves(seed)
    lib().orc_csprng_fill_u64(_p(out), C.c_uint64(first_word), C.c_uint64(count), lo, hi)
    return out


def decompress_seeded_bsk(seeded, n, k, N, level, seed: int) -> np.ndarray:
    assert seeded.size == n * level * (k + 1) * N
    out = np.empty(n * level * (k + 1) ** 2 * N, dtype=np.uint64)
    lo, hi = _seed_halves(seed)
    lib().orc_decompress_seeded_bsk(_p(out), _p(np.ascontiguousarray(seeded)), n, k, C.c_size_t(N),
                                    level, lo, hi)
    return out


def decompress_seeded_ksk(seeded, n_in, n_out, level, seed: int) -> np.ndarray:
    assert seeded.size == n_in * level
    out = np.empty(n_in * level * (n_out + 1), dtype=np.uint64)
    lo, hi = _seed_halves(seed)
    lib().orc_decompress_seeded_ksk(_p(out), _p(np.ascontiguousarray(seeded)), n_in, n_out, level,
                                    lo, hi)
    return out


# ---------------------------------------------------------------- WoP-PBS (wop_oracle.inl)
def gen_cbs_pfpksk_list(lwe_key_in, glwe_key, k, N, base_log, level, sigma, seed) -> np.ndarray:
    """(k+1) private functional packing keyswitch keys [k+1][n_in+1][level][(k+1)N]."""
    n_in = len(lwe_key_in)
    out = np.empty((k + 1) * (n_in + 1) * level * (k + 1) * N, dtype=np.uint64)
    lib().orc_gen_cbs_pfpksk_list(_p(out), _p(lwe_key_in), n_in, _p(glwe_key), k, C.c_size_t(N),
                                  base_log, level, C.c_double(sigma), C.c_uint64(seed))
    return out


def extract_bits(lwe_in, fbsk, ksk, n_small, k, N, bsk_b, bsk_l, ks_b, ks_l, delta_log, nbits):
    out = np.empty((nbits, n_small + 1), dtype=np.uint64)
    lib().orc_extract_bits(_p(out), _p(np.ascontiguousarray(lwe_in)), _p(fbsk, np.complex128), _p(ksk),
                           n_small, k, C.c_size_t(N), bsk_b, bsk_l, ks_b, ks_l, delta_log, nbits)
    return out


def circuit_bootstrap_boolean(lwe_in, fbsk, pfpksk, n_small, k, N, bsk_b, bsk_l, pks_b, pks_l,
                              cbs_b, cbs_l, delta_log=63):
    out = np.empty(cbs_l * (k + 1) * (k + 1) * N, dtype=np.uint64)
    lib().orc_circuit_bootstrap_boolean(_p(out), _p(np.ascontiguousarray(lwe_in)),
                                        _p(fbsk, np.complex128), _p(pfpksk), n_small, k,
                                        C.c_size_t(N), bsk_b, bsk_l, pks_b, pks_l, cbs_b, cbs_l,
                                        delta_log)
    return out


def cbs_vertical_packing(bits_in, luts, fbsk, pfpksk, n_small, k, N, bsk_b, bsk_l, pks_b, pks_l,
                         cbs_b, cbs_l):
    nbits = bits_in.shape[0]
    lut_count, lut_size = luts.shape
    out = np.empty((lut_count, k * N + 1), dtype=np.uint64)
    lib().orc_cbs_vertical_packing(_p(out), _p(np.ascontiguousarray(bits_in)),
                                   _p(np.ascontiguousarray(luts)), _p(fbsk, np.complex128),
                                   _p(pfpksk), n_small, k, C.c_size_t(N), bsk_b, bsk_l, pks_b, pks_l,
                                   cbs_b, cbs_l, nbits, C.c_size_t(lut_count), C.c_size_t(lut_size))
    return out


def wop_pbs_crt(cts_in, lut_ct, crt_decomp, fbsk, ksk, pfpksk, n_small, k, N, bsk_b, bsk_l, ks_b,
                ks_l, pks_b, pks_l, cbs_b, cbs_l):
    blocks = cts_in.shape[0]
    out = np.empty_like(cts_in)
    lib().orc_wop_pbs_crt(_p(out), _p(np.ascontiguousarray(cts_in)), _p(np.ascontiguousarray(lut_ct)),
                          _p(np.ascontiguousarray(crt_decomp, dtype=np.uint64)), blocks,
                          _p(fbsk, np.complex128), _p(ksk), _p(pfpksk), n_small, k, C.c_size_t(N),
                          bsk_b, bsk_l, ks_b, ks_l, pks_b, pks_l, cbs_b, cbs_l)
    return out
