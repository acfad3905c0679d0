"""ctypes binding of the REFERENCE GPU library (vendored tfhe-cuda-backend) built by
oracle/ref_build.mk into oracle/_ref/libtfhe_cuda_backend_ref.so.

TEST / BENCH INFRASTRUCTURE ONLY (same rule as the rest of oracle/): a second parity witness and
"the kernel to beat" timed on the same box.  Same entry-point names as our library, so it is
loaded RTLD_LOCAL from its own handle (both libraries are linked -Bsymbolic).
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
PATH = os.path.join(_HERE, "_ref", "libtfhe_cuda_backend_ref.so")
_LIB = None

vp, u32, i8p = C.c_void_p, C.c_uint32, C.POINTER(C.c_int8)
i8pp = C.POINTER(i8p)


def available() -> bool:
    return os.path.exists(PATH)


def lib():
    global _LIB
    if _LIB is not None:
        return _LIB
    l = C.CDLL(PATH, mode=C.RTLD_LOCAL)
    sig = {
        "cuda_keyswitch_lwe_ciphertext_vector_64": [vp, u32, vp, vp, vp, vp, vp, u32, u32, u32, u32, u32],
        "cuda_convert_lwe_programmable_bootstrap_key_64": [vp, u32, vp, vp, u32, u32, u32, u32],
        "scratch_cuda_programmable_bootstrap_64": [vp, u32, i8pp, u32, u32, u32, u32, C.c_bool],
        "cuda_programmable_bootstrap_lwe_ciphertext_vector_64":
            [vp, u32, vp, vp, vp, vp, vp, vp, vp, i8p, u32, u32, u32, u32, u32, u32, u32, u32],
        "cleanup_cuda_programmable_bootstrap": [vp, u32, i8pp],
        "cuda_synchronize_stream": [vp, u32],
    }
    for name, args in sig.items():
        f = getattr(l, name)
        f.restype = None
        f.argtypes = args
    _LIB = l
    return l


class RefGpu:
    """Drives the reference library with torch tensors as device buffers (stream = torch's)."""

    def __init__(self, device: int = 0):
        import torch

        self.torch = torch
        self.lib = lib()
        self.gpu = device
        self.dev = torch.device("cuda", device)
        self._scratch = {}

    @property
    def stream(self):
        return self.torch.cuda.current_stream(self.dev).cuda_stream

    def convert_bsk(self, bsk_host, n, k, level, N):
        torch = self.torch
        dest = torch.empty(bsk_host.size, dtype=torch.int64, device=self.dev)
        self.lib.cuda_convert_lwe_programmable_bootstrap_key_64(
            self.stream, self.gpu, dest.data_ptr(), bsk_host.ctypes.data, n, k, level, N)
        self.lib.cuda_synchronize_stream(self.stream, self.gpu)
        return dest

    def keyswitch(self, cts, ksk, n_in, n_out, base_log, level, out=None):
        torch = self.torch
        num = int(cts.shape[0])
        if out is None:
            out = torch.empty((num, n_out + 1), dtype=torch.int64, device=self.dev)
        idx = torch.arange(num, dtype=torch.int64, device=self.dev)
        self.lib.cuda_keyswitch_lwe_ciphertext_vector_64(
            self.stream, self.gpu, out.data_ptr(), idx.data_ptr(), cts.data_ptr(), idx.data_ptr(),
            ksk.data_ptr(), n_in, n_out, base_log, level, num)
        return out

    def scratch(self, k, N, level, count):
        key = (k, N, level, count)
        if key not in self._scratch:
            buf = i8p()
            self.lib.scratch_cuda_programmable_bootstrap_64(self.stream, self.gpu, C.byref(buf), k,
                                                            N, level, count, True)
            self._scratch[key] = buf
        return self._scratch[key]

    def bootstrap(self, cts, luts, fbsk, n, k, N, base_log, level, out=None):
        torch = self.torch
        num = int(cts.shape[0])
        if out is None:
            out = torch.empty((num, k * N + 1), dtype=torch.int64, device=self.dev)
        idx = torch.arange(num, dtype=torch.int64, device=self.dev)
        lut_idx = torch.zeros(num, dtype=torch.int64, device=self.dev)
        buf = self.scratch(k, N, level, num)
        self.lib.cuda_programmable_bootstrap_lwe_ciphertext_vector_64(
            self.stream, self.gpu, out.data_ptr(), idx.data_ptr(), luts.data_ptr(),
            lut_idx.data_ptr(), cts.data_ptr(), idx.data_ptr(), fbsk.data_ptr(), buf, n, k, N,
            base_log, level, num, 1, 1)
        return out

    def close(self):
        for buf in self._scratch.values():
            self.lib.cleanup_cuda_programmable_bootstrap(self.stream, self.gpu, C.byref(buf))
        self._scratch.clear()
