"""Parameter sets of the workloads in BASELINE.json / SURVEY.md §8.

Sources (under /root/reference): compilers/concrete-optimizer/v0-parameters/ref/v0_last_128 and
third_party/tfhe-rs/tfhe/src/core_crypto/algorithms/test/mod.rs:56-78 (P2048).
Noise: 128-bit security curve of tools/parameter-curves/concrete-security-curves-rust/
src/gaussian/{curves_gen.rs:2-11,security.rs:30-45}: log2(sigma) = max(a*d + b, -62).
"""
from __future__ import annotations

from dataclasses import dataclass


def security_sigma(dim: int) -> float:
    a, b = -0.025696778711484593, 2.675931372549016
    return 2.0 ** max(a * dim + b, 2 - 64)


@dataclass(frozen=True)
class TfheParams:
    name: str
    n: int          # small LWE dimension
    k: int          # GLWE dimension
    N: int          # polynomial size
    l: int          # PBS levels
    b: int          # PBS base log
    l_ks: int       # keyswitch levels
    b_ks: int       # keyswitch base log
    msg_bits: int   # message precision the set was optimised for
    sigma_lwe: float = 0.0
    sigma_glwe: float = 0.0

    @property
    def big_n(self) -> int:
        return self.k * self.N

    @property
    def bsk_bytes(self) -> int:
        return self.n * (self.k + 1) ** 2 * self.l * self.N * 8

    @property
    def ksk_bytes(self) -> int:
        return self.big_n * self.l_ks * (self.n + 1) * 8

    def with_noise(self) -> "TfheParams":
        if self.sigma_lwe and self.sigma_glwe:
            return self
        return TfheParams(**{**self.__dict__, "sigma_lwe": security_sigma(self.n),
                             "sigma_glwe": security_sigma(self.big_n)})

    # algorithmic counts of SURVEY.md §8(d) / BASELINE.md
    def fft_flops(self) -> float:
        import math
        h = self.N // 2
        return 5 * h * math.log2(h) + 6 * h

    def pbs_flops(self) -> float:
        k, l, h = self.k, self.l, self.N // 2
        return self.n * ((k + 1) * (l + 1) * self.fft_flops() + (k + 1) ** 2 * l * h * 8)

    def pbs_bytes(self, samples: int, n_luts: int = 1) -> float:
        return self.bsk_bytes + samples * 8 * ((self.n + 1) + (self.big_n + 1)) + n_luts * (self.k + 1) * self.N * 8

    def ks_macs(self) -> float:
        return self.big_n * self.l_ks * (self.n + 1)

    def ks_bytes(self, samples: int) -> float:
        return self.ksk_bytes + samples * 8 * ((self.big_n + 1) + (self.n + 1))


PARAMS = {
    # cfg 1 / cfg 5 (4-bit)
    "P4": TfheParams("P4", 801, 2, 1024, 1, 23, 3, 4, 4).with_noise(),
    # cfg 2 microbench (tfhe-rs TEST_PARAMS_4_BITS_NATIVE_U64 / benchmark_pbs.cpp:395)
    "P2048": TfheParams("P2048", 742, 1, 2048, 1, 23, 5, 3, 4, 7.06984e-6, 2.94036e-16),
    "P6": TfheParams("P6", 880, 1, 4096, 1, 22, 4, 4, 6).with_noise(),
    # cfg 3
    "P7": TfheParams("P7", 915, 1, 8192, 1, 22, 6, 3, 7).with_noise(),
    "P8": TfheParams("P8", 1006, 1, 16384, 2, 15, 5, 4, 8).with_noise(),
    # toy sets for CPU-side tests (not secure)
    "TOY512": TfheParams("TOY512", 32, 1, 512, 2, 10, 4, 4, 3, 2.0 ** -20, 2.0 ** -45),
    "TOY256K2": TfheParams("TOY256K2", 24, 2, 256, 3, 7, 3, 5, 2, 2.0 ** -20, 2.0 ** -45),
    "TOY1024": TfheParams("TOY1024", 40, 1, 1024, 1, 23, 5, 3, 3, 2.0 ** -22, 2.0 ** -50),
}
