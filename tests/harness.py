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

    def __init__(self, p: TfheParams, seed: int = 42, fourier_oracle: bool = True, n_override=None,
                 compression_seeds=None):
        """`compression_seeds` = (ksk_seed, bsk_seed), two u128: generate SEEDED keys — the masks
        are the reference CSPRNG's stream for that seed (oracle/csprng_oracle.cpp) and
        `seeded_ksk` / `seeded_bsk` hold the bodies only, as Concrete ships them
        (generate_seeded_lwe_{keyswitch,bootstrap}_key); `ksk` / `bsk` are the full keys."""
        if n_override is not None:
            p = TfheParams(**{**p.__dict__, "n": n_override, "name": f"{p.name}_n{n_override}"})
        self.p = p
        self.sk_small = orc.gen_binary_key(p.n, seed)
        self.sk_glwe = orc.gen_binary_key(p.k * p.N, seed + 1)  # == big LWE key (into_lwe_secret_key)
        if compression_seeds is None:
            self.ksk = orc.gen_ksk(self.sk_glwe, self.sk_small, p.b_ks, p.l_ks, p.sigma_lwe, seed + 2)
            self.bsk = orc.gen_bsk(self.sk_small, self.sk_glwe, p.k, p.N, p.b, p.l, p.sigma_glwe, seed + 3)
        else:
            self.ksk_seed, self.bsk_seed = compression_seeds
            km = orc.csprng_u64(self.ksk_seed, 0, p.big_n * p.l_ks * p.n)
            self.ksk = orc.gen_ksk_masks(self.sk_glwe, self.sk_small, p.b_ks, p.l_ks, p.sigma_lwe,
                                         seed + 2, km)
            bm = orc.csprng_u64(self.bsk_seed, 0, p.n * p.l * (p.k + 1) * p.k * p.N)
            self.bsk = orc.gen_bsk_masks(self.sk_small, self.sk_glwe, p.k, p.N, p.b, p.l,
                                         p.sigma_glwe, seed + 3, bm)
            self.seeded_ksk = np.ascontiguousarray(
                self.ksk.reshape(p.big_n * p.l_ks, p.n + 1)[:, p.n])
            self.seeded_bsk = np.ascontiguousarray(
                self.bsk.reshape(p.n * p.l * (p.k + 1), p.k + 1, p.N)[:, p.k]).reshape(-1)
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
    with np.errstate(over="ignore"):  # wrapping negation on purpose
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


def phase_distance(ks: "KeySet", a: np.ndarray, b: np.ndarray) -> np.ndarray:
    """|phase(a) - phase(b)| per ciphertext under the big key, in units of 2^-64.

    Why phases and not ciphertext words: from the second CMUX on, an f64 rounding difference of
    a few ulp in the accumulator flips the signed digit of any coefficient that sits next to a
    rounding boundary of closest_representable, and one flipped digit adds +-1 x (a uniformly
    random BSK mask row) to the output mask.  Ciphertext WORDS of two f64 implementations (or of
    concrete-cpu on two machines) therefore differ by ~2^63 while the PHASE
    (b - <a, s>, i.e. plaintext + noise) stays within the PBS noise (see phase_tolerance).  The
    word-level tolerance is checked where it is meaningful: one external product from an
    identical accumulator (tests/test_gpu_parity.py::test_external_product_word_tolerance)."""
    pa = orc.lwe_decrypt(ks.sk_glwe, np.ascontiguousarray(a))
    pb = orc.lwe_decrypt(ks.sk_glwe, np.ascontiguousarray(b))
    return torus_distance(pa, pb)


def sigma_round_ulp(p: TfheParams) -> float:
    """Std of the decomposition-rounding noise a full PBS adds to the phase, units of 2^-64:
    every ct1 coefficient is rounded to a multiple of the granule g = 2^(64 - b*l) (error uniform
    in +-g/2, variance g^2/12) and reaches the phase through <mask, key> (k*N/2 + 1 terms), once
    per CMUX (noise model: backends/concrete-cpu/noise-model/src/gaussian_noise/noise/
    external_product_glwe.rs:62-89, first term)."""
    import math
    g = 2.0 ** (64 - p.b * p.l)
    return math.sqrt(p.n * (p.k * p.N / 2 + 1) / 12.0) * g


def phase_tolerance(p: TfheParams) -> float:
    """Gate for |phase(ours) - phase(oracle)| after a FULL PBS, in units of 2^-64.

    Two f64 implementations agree word for word only until the first signed digit flips at a
    rounding boundary of closest_representable (an accumulator difference of a few ulp is
    enough).  From that CMUX on the two runs hold different — equally valid — roundings of their
    accumulators, i.e. independent realisations of the PBS's own decomposition-rounding noise, so
    their phases differ like two noise samples: std sqrt(2) * sigma_round on top of the FFT
    error.  Gate = 8 * sqrt(sigma_fft^2 + sigma_round^2).  P2048: 2^52.5, i.e. 1% of a 4-bit
    message step (2^59).  The tight word-level gate (8 * sigma_fft) is applied where it is
    meaningful: one CMUX from an identical accumulator."""
    import math
    return 8.0 * math.sqrt(sigma_fft_ulp(p) ** 2 + sigma_round_ulp(p) ** 2)
