"""Shared test/bench harness: keys, encodings, LUTs (oracle-side; test infrastructure)."""
from __future__ import annotations

import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import oracle as orc  # noqa: E402
from concrete_b200.params import PARAMS, TfheParams  # noqa: E402


class KeySet:
    """One big (GLWE-derived) key, one small key, one KSK, one BSK — all seeded."""

    def __init__(self, p: TfheParams, seed: int = 42, fourier_oracle: bool = True, n_override=None):
        if n_override is not None:
            p = TfheParams(**{**p.__dict__, "n": n_override, "name": f"{p.name}_n{n_override}"})
        self.p = p
        self.sk_small = orc.gen_binary_key(p.n, seed)
        self.sk_glwe = orc.gen_binary_key(p.k * p.N, seed + 1)  # == big LWE key (into_lwe_secret_key)
        self.ksk = orc.gen_ksk(self.sk_glwe, self.sk_small, p.b_ks, p.l_ks, p.sigma_lwe, seed + 2)
        self.bsk = orc.gen_bsk(self.sk_small, self.sk_glwe, p.k, p.N, p.b, p.l, p.sigma_glwe, seed + 3)
        self.fbsk = orc.bsk_to_fourier(self.bsk, p.n, p.k, p.l, p.N) if fourier_oracle else None

    def encrypt_big(self, msgs: np.ndarray, bits: int, seed: int = 7) -> np.ndarray:
        return orc.lwe_encrypt(self.sk_glwe, encode(msgs, bits), self.p.sigma_glwe, seed)

    def encrypt_small(self, msgs: np.ndarray, bits: int, seed: int = 9) -> np.ndarray:
        return orc.lwe_encrypt(self.sk_small, encode(msgs, bits), self.p.sigma_lwe, seed)

    def decrypt_big(self, cts: np.ndarray, bits: int) -> np.ndarray:
        return decode(orc.lwe_decrypt(self.sk_glwe, cts), bits)

    def decrypt_small(self, cts: np.ndarray, bits: int) -> np.ndarray:
        return decode(orc.lwe_decrypt(self.sk_small, cts), bits)


def encode(msgs: np.ndarray, bits: int) -> np.ndarray:
    """m * 2^(63-bits): one padding bit, as memref_encode_expand_lut_for_bootstrap assumes."""
    return (np.asarray(msgs, dtype=np.uint64) << np.uint64(63 - bits)).astype(np.uint64)


def decode(plaintexts: np.ndarray, bits: int) -> np.ndarray:
    shift = np.uint64(63 - bits)
    half = np.uint64(1) << (shift - np.uint64(1))
    return ((plaintexts + half) >> shift) % np.uint64(1 << (bits + 1))


def expand_lut(table: np.ndarray, N: int, out_bits: int) -> np.ndarray:
    """memref_encode_expand_lut_for_bootstrap (compilers/concrete-compiler/compiler/lib/Runtime/
    wrappers.cpp:388-450), unsigned case."""
    table = np.asarray(table, dtype=np.uint64)
    size = len(table)
    mega = N // size
    assert mega % 2 == 0
    shift = np.uint64(64 - out_bits - 1)
    out = np.zeros(N, dtype=np.uint64)
    out[: mega // 2] = table[0] << shift
    out[(size - 1) * mega + mega // 2:] = (np.uint64(0) - (table[0] << shift))
    for i in range(1, size):
        s = mega * (i - 1) + mega // 2
        out[s: s + mega] = table[i] << shift
    return out


def trivial_glwe(lut_poly: np.ndarray, k: int) -> np.ndarray:
    """Trivial GLWE accumulator (wrappers.cpp:199-213): zero mask polys, body = LUT."""
    N = len(lut_poly)
    g = np.zeros((k + 1) * N, dtype=np.uint64)
    g[k * N:] = lut_poly
    return g


def torus_distance(a: np.ndarray, b: np.ndarray) -> np.ndarray:
    """|a - b| on the torus in units of 2^-64, as float64."""
    d = (a - b).astype(np.uint64)
    s = d.view(np.int64)
    return np.abs(s.astype(np.float64))


def sigma_fft_ulp(p: TfheParams) -> float:
    """f64-FFT error budget of one PBS in units of 2^-64 (noise-model
    external_product_glwe.rs:62-89 summed over n iterations)."""
    import math
    return math.sqrt(p.n * 2 ** 19.42 * p.l * (p.k + 1)) * (2 ** p.b) * p.N
