"""Host-side mirror of the reference's GPU call sequence, one object per device.

`Engine` plays the role of Concrete's RuntimeContext + GPU wrappers
(compilers/concrete-compiler/compiler/include/concretelang/Runtime/context.h:86-145 and
lib/Runtime/wrappers.cpp:110-256): it keeps the per-device key copies, the PBS scratch and the
index arrays alive between calls, and drives the C ABI with raw device pointers.  torch is used
only to own device memory and the stream; every compute call goes through libconcrete_b200.so.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from .lib import load_library


def as_i64(a: np.ndarray) -> torch.Tensor:
    """numpy uint64 -> torch int64 view (same bits)."""
    return torch.from_numpy(np.ascontiguousarray(a).view(np.int64))


def as_u64(t: torch.Tensor) -> np.ndarray:
    """torch int64 (any device) -> numpy uint64 (same bits)."""
    return t.detach().cpu().contiguous().numpy().view(np.uint64)


class Engine:
    def __init__(self, device: int = 0):
        if not torch.cuda.is_available():
            raise RuntimeError("concrete_b200.Engine needs a CUDA device (no CPU fallback)")
        self.lib = load_library()
        self.gpu = int(device)
        self.device = torch.device("cuda", self.gpu)
        torch.cuda.set_device(self.device)
        self._iota = None
        self._scratch = {}  # (k, N, l, count) -> int8*

    # ------------------------------------------------------------------ plumbing
    @property
    def stream(self) -> int:
        return torch.cuda.current_stream(self.device).cuda_stream

    def to_device(self, a: np.ndarray) -> torch.Tensor:
        return as_i64(a).to(self.device, non_blocking=False)

    def iota(self, n: int) -> torch.Tensor:
        if self._iota is None or self._iota.numel() < n:
            self._iota = torch.arange(max(n, 1024), dtype=torch.int64, device=self.device)
        return self._iota

    def synchronize(self):
        self.lib.cuda_synchronize_device(self.gpu)

    # ------------------------------------------------------------------ keys
    def upload_ksk(self, ksk: np.ndarray) -> torch.Tensor:
        """Raw H2D copy, like RuntimeContext::get_ksk_gpu (context.h:117-145)."""
        return self.to_device(ksk)

    def convert_bsk(self, bsk: np.ndarray, n: int, k: int, level: int, N: int) -> torch.Tensor:
        """cuda_convert_lwe_programmable_bootstrap_key_64 from a host standard-domain key,
        like RuntimeContext::get_bsk_gpu (context.h:86-115)."""
        bsk = np.ascontiguousarray(bsk, dtype=np.uint64)
        assert bsk.size == n * level * (k + 1) ** 2 * N
        dest = torch.empty(bsk.size, dtype=torch.int64, device=self.device)
        self.lib.cuda_convert_lwe_programmable_bootstrap_key_64(
            self.stream, self.gpu, dest.data_ptr(), bsk.ctypes.data, n, k, level, N)
        self.lib.cuda_synchronize_stream(self.stream, self.gpu)  # as context.h:110
        return dest

    # ------------------------------------------------------------------ seeded keys
    @staticmethod
    def _halves(seed: int):
        return seed & (2 ** 64 - 1), seed >> 64

    def decompress_seeded_bsk(self, seeded: np.ndarray, n, k, level, N, seed: int,
                              convert: bool = False) -> torch.Tensor:
        """Seeded BSK (bodies, host) -> standard BSK on the device, or (convert=True) straight to
        the Fourier key the PBS consumes; masks regenerated on the device from the 128-bit seed
        (replaces Keys.cpp:193-213 + context.h:86-115)."""
        seeded = np.ascontiguousarray(seeded, dtype=np.uint64)
        assert seeded.size == n * level * (k + 1) * N
        dest = torch.empty(n * level * (k + 1) ** 2 * N, dtype=torch.int64, device=self.device)
        lo, hi = self._halves(seed)
        f = (self.lib.cuda_convert_seeded_lwe_programmable_bootstrap_key_64 if convert
             else self.lib.cuda_decompress_seeded_lwe_bootstrap_key_64)
        f(self.stream, self.gpu, dest.data_ptr(), seeded.ctypes.data, n, k, level, N, lo, hi)
        self.lib.cuda_synchronize_stream(self.stream, self.gpu)  # host bodies may go away now
        return dest

    def decompress_seeded_ksk(self, seeded: np.ndarray, n_in, n_out, level, seed: int) -> torch.Tensor:
        seeded = np.ascontiguousarray(seeded, dtype=np.uint64)
        assert seeded.size == n_in * level
        dest = torch.empty(n_in * level * (n_out + 1), dtype=torch.int64, device=self.device)
        lo, hi = self._halves(seed)
        self.lib.cuda_decompress_seeded_lwe_keyswitch_key_64(
            self.stream, self.gpu, dest.data_ptr(), seeded.ctypes.data, n_in, n_out, level, lo, hi)
        self.lib.cuda_synchronize_stream(self.stream, self.gpu)
        return dest

    # ------------------------------------------------------------------ keyswitch
    def keyswitch(self, cts: torch.Tensor, ksk: torch.Tensor, n_in: int, n_out: int,
                  base_log: int, level: int, out: torch.Tensor | None = None,
                  in_idx: torch.Tensor | None = None, out_idx: torch.Tensor | None = None,
                  num_samples: int | None = None) -> torch.Tensor:
        num = int(cts.shape[0]) if num_samples is None else num_samples
        if out is None:
            out = torch.empty((num, n_out + 1), dtype=torch.int64, device=self.device)
        in_idx = self.iota(num) if in_idx is None else in_idx
        out_idx = self.iota(num) if out_idx is None else out_idx
        self.lib.cuda_keyswitch_lwe_ciphertext_vector_64(
            self.stream, self.gpu, out.data_ptr(), out_idx.data_ptr(), cts.data_ptr(),
            in_idx.data_ptr(), ksk.data_ptr(), n_in, n_out, base_log, level, num)
        return out

    # ------------------------------------------------------------------ PBS
    def pbs_scratch(self, k: int, N: int, level: int, count: int):
        key = (k, N, level)
        ent = self._scratch.get(key)
        if ent is not None and ent[1] >= count:
            return ent[0]
        if ent is not None:
            self._cleanup(ent[0])
        buf = C.POINTER(C.c_int8)()
        self.lib.scratch_cuda_programmable_bootstrap_64(
            self.stream, self.gpu, C.byref(buf), k, N, level, count, True)
        self._scratch[key] = (buf, count)
        return buf

    def _cleanup(self, buf):
        self.lib.cleanup_cuda_programmable_bootstrap(self.stream, self.gpu, C.byref(buf))

    def close(self):
        for buf, _ in self._scratch.values():
            self._cleanup(buf)
        self._scratch.clear()

    def bootstrap(self, cts: torch.Tensor, luts: torch.Tensor, lut_idx: torch.Tensor | None,
                  fbsk: torch.Tensor, n: int, k: int, N: int, base_log: int, level: int,
                  out: torch.Tensor | None = None, in_idx=None, out_idx=None,
                  lut_count: int = 1, lut_stride: int = 1,
                  num_samples: int | None = None) -> torch.Tensor:
        num = int(cts.shape[0]) if num_samples is None else num_samples
        if out is None:
            out = torch.empty((lut_count * num, k * N + 1), dtype=torch.int64, device=self.device)
        if lut_idx is None:
            lut_idx = torch.zeros(num, dtype=torch.int64, device=self.device)
        in_idx = self.iota(num) if in_idx is None else in_idx
        out_idx = self.iota(num) if out_idx is None else out_idx
        buf = self.pbs_scratch(k, N, level, num)
        self.lib.cuda_programmable_bootstrap_lwe_ciphertext_vector_64(
            self.stream, self.gpu, out.data_ptr(), out_idx.data_ptr(), luts.data_ptr(),
            lut_idx.data_ptr(), cts.data_ptr(), in_idx.data_ptr(), fbsk.data_ptr(), buf,
            n, k, N, base_log, level, num, lut_count, lut_stride)
        return out

    def keyswitch_bootstrap(self, cts, ksk, fbsk, luts, lut_idx, p, out=None, ks_tmp=None):
        """KS -> PBS chained on the device (what GPUDFG.cpp:1057-1250 does per chunk)."""
        num = int(cts.shape[0])
        if out is None:
            out = torch.empty((num, p.big_n + 1), dtype=torch.int64, device=self.device)
        if ks_tmp is None:
            ks_tmp = torch.empty((num, p.n + 1), dtype=torch.int64, device=self.device)
        if lut_idx is None:
            lut_idx = torch.zeros(num, dtype=torch.int64, device=self.device)
        buf = self.pbs_scratch(p.k, p.N, p.l, num)
        self.lib.cuda_keyswitch_bootstrap_lwe_ciphertext_vector_64(
            self.stream, self.gpu, out.data_ptr(), luts.data_ptr(), lut_idx.data_ptr(),
            cts.data_ptr(), ksk.data_ptr(), fbsk.data_ptr(), buf, ks_tmp.data_ptr(),
            p.big_n, p.n, p.b_ks, p.l_ks, p.k, p.N, p.b, p.l, num)
        return out

    # ------------------------------------------------------------------ WoP-PBS
    def extract_bits(self, cts: torch.Tensor, nbits, fbsk, ksk, p) -> torch.Tensor:
        """`p`: an object with n, k, N, bsk_b, bsk_l, ks_b, ks_l (tests/wop_harness.WopParams)."""
        nb = (C.c_uint32 * len(nbits))(*nbits)
        out = torch.zeros((sum(nbits), p.n + 1), dtype=torch.int64, device=self.device)
        self.lib.cuda_extract_bits_lwe_ciphertext_vector_64(
            self.stream, self.gpu, out.data_ptr(), cts.data_ptr(), nb, len(nbits), fbsk.data_ptr(),
            ksk.data_ptr(), p.n, p.k, p.N, p.bsk_l, p.bsk_b, p.ks_l, p.ks_b)
        return out

    def circuit_bootstrap_vertical_packing(self, bits: torch.Tensor, luts: torch.Tensor, fbsk, pfpksk,
                                           p) -> torch.Tensor:
        nbits, lut_count, lut_size = int(bits.shape[0]), int(luts.shape[0]), int(luts.shape[1])
        out = torch.empty((lut_count, p.k * p.N + 1), dtype=torch.int64, device=self.device)
        self.lib.cuda_circuit_bootstrap_vertical_packing_lwe_ciphertext_vector_64(
            self.stream, self.gpu, out.data_ptr(), bits.data_ptr(), luts.data_ptr(), fbsk.data_ptr(),
            pfpksk.data_ptr(), nbits, lut_count, lut_size, p.n, p.k, p.N, p.bsk_l, p.bsk_b, p.pks_l,
            p.pks_b, p.cbs_l, p.cbs_b)
        return out

    def wop_pbs_crt(self, cts: torch.Tensor, lut_ct: torch.Tensor, moduli, fbsk, ksk, pfpksk,
                    p) -> torch.Tensor:
        """memref_wop_pbs_crt_buffer (wrappers.cpp:855-997) with device-resident operands."""
        crt = (C.c_uint64 * len(moduli))(*moduli)
        out = torch.empty_like(cts)
        self.lib.cuda_wop_pbs_crt_lwe_ciphertext_vector_64(
            self.stream, self.gpu, out.data_ptr(), cts.data_ptr(), lut_ct.data_ptr(), crt,
            len(moduli), fbsk.data_ptr(), ksk.data_ptr(), pfpksk.data_ptr(), p.n, p.k, p.N, p.bsk_l,
            p.bsk_b, p.ks_l, p.ks_b, p.pks_l, p.pks_b, p.cbs_l, p.cbs_b)
        return out

    def fourier_polynomial_mul(self, a: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
        total, N = a.shape
        out = torch.empty_like(a)
        self.lib.cuda_fourier_polynomial_mul(self.stream, self.gpu, a.data_ptr(), b.data_ptr(),
                                             out.data_ptr(), N, total)
        return out
