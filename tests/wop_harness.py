"""Harness for the WoP-PBS tests (test infrastructure): parameter sets, keys, and the runtime's CRT
encodings restated from compilers/concrete-compiler/compiler/lib/Common/CRT.cpp:80-95 and
lib/Runtime/wrappers.cpp:452-575 (memref_encode_lut_for_crt_woppbs, unsigned case)."""
from __future__ import annotations

import math
from dataclasses import dataclass

import numpy as np

from harness import orc


@dataclass(frozen=True)
class WopParams:
    name: str
    n: int        # small LWE dimension
    k: int
    N: int
    bsk_b: int
    bsk_l: int
    ks_b: int
    ks_l: int
    pks_b: int
    pks_l: int
    cbs_b: int
    cbs_l: int
    sigma_lwe: float = 2.0 ** -40
    sigma_glwe: float = 2.0 ** -52

    @property
    def big_n(self) -> int:
        return self.k * self.N


WOP_PARAMS = {
    # lut of 2^9 entries over N = 256: one cmux-tree layer + 8 blind-rotation steps
    "WOP256": WopParams("WOP256", 48, 1, 256, 7, 4, 4, 5, 8, 4, 6, 3),
    # k = 2: three private functional keyswitch keys, GGSWs of 3 x 3 polynomials
    "WOP256K2": WopParams("WOP256K2", 40, 2, 256, 7, 4, 4, 5, 8, 4, 6, 3),
    # the shape of the reference's FFT_WOPBS_PARAMS (tfhe core_crypto/algorithms/test/mod.rs:350-372:
    # k = 1, N = 1024, wide bases split over many levels), small dimension cut for the oracle
    "WOP1024": WopParams("WOP1024", 64, 1, 1024, 4, 9, 1, 9, 4, 9, 6, 4),
    # the optimizer's parameters for 16 bits of precision through WoP-PBS at 128-bit security
    # (compilers/concrete-optimizer/v0-parameters/ref/wop_pbs_last_128, "16: # bits", row 0:
    # k=2, N=2^10, n=707, br_l=4, br_b=9, ks_l=6, ks_b=2, cb_l=2, cb_b=9, pp_l=2, pp_b=16), noise
    # from the 128-bit security curve like every other full-size set
    "WOP16": WopParams("WOP16", 707, 2, 1024, 9, 4, 2, 6, 16, 2, 9, 2,
                       sigma_lwe=2.0 ** max(-0.025696778711484593 * 707 + 2.675931372549016, -62),
                       sigma_glwe=2.0 ** max(-0.025696778711484593 * 2048 + 2.675931372549016, -62)),
}


class WopKeySet:
    def __init__(self, p: WopParams, seed: int = 42):
        self.p = p
        self.sk_small = orc.gen_binary_key(p.n, seed)
        self.sk_glwe = orc.gen_binary_key(p.k * p.N, seed + 1)
        self.ksk = orc.gen_ksk(self.sk_glwe, self.sk_small, p.ks_b, p.ks_l, p.sigma_lwe, seed + 2)
        self.bsk = orc.gen_bsk(self.sk_small, self.sk_glwe, p.k, p.N, p.bsk_b, p.bsk_l, p.sigma_glwe,
                               seed + 3)
        self.fbsk = orc.bsk_to_fourier(self.bsk, p.n, p.k, p.bsk_l, p.N)
        self.pfpksk = orc.gen_cbs_pfpksk_list(self.sk_glwe, self.sk_glwe, p.k, p.N, p.pks_b, p.pks_l,
                                              p.sigma_glwe, seed + 4)

    def encrypt_big_raw(self, plaintexts, seed=7):
        return orc.lwe_encrypt(self.sk_glwe, np.asarray(plaintexts, dtype=np.uint64),
                               self.p.sigma_glwe, seed)

    def decrypt_big_raw(self, cts):
        return orc.lwe_decrypt(self.sk_glwe, np.ascontiguousarray(cts))

    def decrypt_small_raw(self, cts):
        return orc.lwe_decrypt(self.sk_small, np.ascontiguousarray(cts))


def crt_bits(moduli):
    return [int(math.ceil(math.log2(m))) for m in moduli]


def crt_encode(value: int, modulus: int, product: int) -> int:
    """concretelang::crt::encode (CRT.cpp:80-88)."""
    if value < 0:
        value += product
    return ((value % modulus) << 64) // modulus & (2 ** 64 - 1)


def crt_decode(val: int, modulus: int) -> int:
    """concretelang::crt::decode (CRT.cpp:90-95)."""
    r = int(val) * modulus
    r += (r & (1 << 63)) << 1
    return (r >> 64) % modulus


def encode_lut_for_crt_woppbs(table, moduli):
    """memref_encode_lut_for_crt_woppbs, unsigned (wrappers.cpp:452-575): one expanded LUT of
    2^total_bits entries per output CRT block, indexed by the concatenated bit fields."""
    bits = crt_bits(moduli)
    product = int(np.prod(moduli))
    table = np.asarray(table, dtype=np.int64)
    assert product >= len(table)
    size = 1 << sum(bits)
    out = np.zeros((len(moduli), size), dtype=np.uint64)
    idx = np.arange(len(table), dtype=np.int64)
    out_index, total = np.zeros(len(table), dtype=np.int64), 0
    for base, nb in zip(moduli, bits):
        out_index += (((idx % base) << nb) // base) << total
        total += nb
    for blk, base in enumerate(moduli):
        # crt::encode (CRT.cpp:80-88): (value mod base) * 2^64 / base, exact in Python integers
        enc = np.array([((int(v) % base) << 64) // base for v in range(base)], dtype=np.uint64)
        out[blk, out_index] = enc[(table % product) % base]
    return out
