"""Python face of the runtime wrapper layer (include/concrete_b200_runtime.h).

`Runtime` owns a RuntimeContext (csrc/runtime_context.h) and calls the `memref_batched_*`
wrappers with HOST numpy buffers exactly as Concrete's generated code would
(compilers/concrete-compiler/compiler/lib/Runtime/wrappers.cpp:110-363): host->device and
device->host copies happen inside the call.
"""
from __future__ import annotations

import numpy as np

from .lib import load_library


def _m2(a: np.ndarray):
    assert a.dtype == np.uint64 and a.ndim == 2 and a.flags["C_CONTIGUOUS"]
    p = a.ctypes.data
    return [p, p, 0, a.shape[0], a.shape[1], a.shape[1], 1]


def _m1(a: np.ndarray):
    assert a.dtype == np.uint64 and a.ndim == 1 and a.flags["C_CONTIGUOUS"]
    p = a.ctypes.data
    return [p, p, 0, a.shape[0], 1]


class Runtime:
    def __init__(self, num_gpus: int = 1, device_base: int = 0):
        self.lib = load_library()
        self.ctx = self.lib.concrete_b200_context_create()
        self.lib.concrete_b200_context_set_device_base(self.ctx, device_base)
        self.lib.concrete_b200_context_set_num_gpus(self.ctx, num_gpus)
        self._keep = []  # host key buffers must outlive the context

    def close(self):
        if self.ctx:
            self.lib.concrete_b200_context_destroy(self.ctx)
            self.ctx = None

    def add_ksk(self, ksk: np.ndarray) -> int:
        ksk = np.ascontiguousarray(ksk, dtype=np.uint64)
        self._keep.append(ksk)
        return int(self.lib.concrete_b200_context_add_ksk(self.ctx, ksk.ctypes.data, ksk.size))

    def add_bsk(self, bsk: np.ndarray) -> int:
        bsk = np.ascontiguousarray(bsk, dtype=np.uint64)
        self._keep.append(bsk)
        return int(self.lib.concrete_b200_context_add_bsk(self.ctx, bsk.ctypes.data, bsk.size))

    def add_seeded_ksk(self, bodies: np.ndarray, full_len: int, seed: int) -> int:
        """Seeded keyswitch key (bodies + u128 compression seed); masks regenerated per device."""
        bodies = np.ascontiguousarray(bodies, dtype=np.uint64)
        self._keep.append(bodies)
        return int(self.lib.concrete_b200_context_add_seeded_ksk(
            self.ctx, bodies.ctypes.data, full_len, seed & (2 ** 64 - 1), seed >> 64))

    def add_seeded_bsk(self, bodies: np.ndarray, full_len: int, seed: int) -> int:
        bodies = np.ascontiguousarray(bodies, dtype=np.uint64)
        self._keep.append(bodies)
        return int(self.lib.concrete_b200_context_add_seeded_bsk(
            self.ctx, bodies.ctypes.data, full_len, seed & (2 ** 64 - 1), seed >> 64))

    def add_pksk(self, pksk: np.ndarray) -> int:
        """The (k+1) private functional packing keyswitch keys of the circuit bootstrap."""
        pksk = np.ascontiguousarray(pksk, dtype=np.uint64)
        self._keep.append(pksk)
        return int(self.lib.concrete_b200_context_add_pksk(self.ctx, pksk.ctypes.data, pksk.size))

    def wop_pbs_crt(self, out, cts, lut_ct, crt_decomp, p, ksk_index=0, bsk_index=0, pksk_index=0):
        """memref_wop_pbs_crt_buffer with HOST buffers (wrappers.cpp:855-997); `p`: WopParams."""
        crt = np.ascontiguousarray(crt_decomp, dtype=np.uint64)
        lut_ct = np.ascontiguousarray(lut_ct, dtype=np.uint64)
        self.lib.memref_wop_pbs_crt_buffer(
            *_m2(out), *_m2(cts), *_m2(lut_ct), *_m1(crt), p.n, p.cbs_l, p.cbs_b, p.ks_l, p.ks_b,
            p.bsk_l, p.bsk_b, p.pks_l, p.pks_b, p.N, ksk_index, bsk_index, pksk_index, self.ctx)
        return out

    def batched_keyswitch(self, out, ct0, level, base_log, n_in, n_out, ksk_index, cpu_named=False):
        f = self.lib.memref_batched_keyswitch_lwe_u64 if cpu_named else self.lib.memref_batched_keyswitch_lwe_cuda_u64
        f(*_m2(out), *_m2(ct0), level, base_log, n_in, n_out, ksk_index, self.ctx)
        return out

    def batched_bootstrap(self, out, ct0, tlu, n, N, level, base_log, k, bsk_index, cpu_named=False):
        f = self.lib.memref_batched_bootstrap_lwe_u64 if cpu_named else self.lib.memref_batched_bootstrap_lwe_cuda_u64
        f(*_m2(out), *_m2(ct0), *_m1(tlu), n, N, level, base_log, k, bsk_index, self.ctx)
        return out

    def batched_mapped_bootstrap(self, out, ct0, tlus, n, N, level, base_log, k, bsk_index):
        self.lib.memref_batched_mapped_bootstrap_lwe_cuda_u64(
            *_m2(out), *_m2(ct0), *_m2(tlus), n, N, level, base_log, k, bsk_index, self.ctx)
        return out

    def keyswitch(self, out, ct0, level, base_log, n_in, n_out, ksk_index):
        self.lib.memref_keyswitch_lwe_cuda_u64(*_m1(out), *_m1(ct0), level, base_log, n_in, n_out,
                                               ksk_index, self.ctx)
        return out

    def bootstrap(self, out, ct0, tlu, n, N, level, base_log, k, bsk_index):
        self.lib.memref_bootstrap_lwe_cuda_u64(*_m1(out), *_m1(ct0), *_m1(tlu), n, N, level,
                                               base_log, k, bsk_index, self.ctx)
        return out

    def batched_keyswitch_bootstrap(self, out, ct0, tlu, p, ksk_index=0, bsk_index=0):
        self.lib.memref_batched_keyswitch_bootstrap_lwe_cuda_u64(
            out.ctypes.data, 0, out.shape[0], out.shape[1], ct0.ctypes.data, 0, ct0.shape[0],
            ct0.shape[1], tlu.ctypes.data, 0, tlu.shape[0], p.l_ks, p.b_ks, p.big_n, p.n,
            ksk_index, p.N, p.l, p.b, p.k, bsk_index, self.ctx)
        return out
