"""Python face of the SDFG ("stream emulator") entry points (include/concrete_b200_runtime.h).

Builds the process graph exactly as Concrete's generated code does under
``fhe.Compiler(..., use_gpu=True)`` with the SDFG lowering
(compilers/concrete-compiler/compiler/lib/Runtime/GPUDFG.cpp:1467-1799): make streams, make
processes connected by them, PUT host memrefs, GET host memrefs.  Used by the tests and by
bench.py's end-to-end leg; every call goes through the C ABI of libconcrete_b200.so.
"""
from __future__ import annotations

import numpy as np

from .lib import load_library

X86_TO_TOPO, TOPO_TO_TOPO, TOPO_TO_X86, TOPO_TO_BOTH, X86_TO_X86 = range(5)


class Graph:
    def __init__(self, runtime, num_gpus: int | None = None):
        """`runtime` is a concrete_b200.runtime.Runtime (owns the RuntimeContext and the keys)."""
        self.lib = load_library()
        self.rt = runtime
        self.dfg = self.lib.stream_emulator_init()
        if num_gpus is not None:
            self.lib.concrete_b200_sdfg_set_num_gpus(self.dfg, num_gpus)
        self._n = 0

    def close(self):
        if self.dfg:
            self.lib.stream_emulator_delete(self.dfg)
            self.dfg = None

    # ---- streams
    def stream(self, stype=TOPO_TO_TOPO, batch=True, name=None):
        self._n += 1
        nm = (name or f"s{self._n}").encode()
        f = self.lib.stream_emulator_make_memref_batch_stream if batch else self.lib.stream_emulator_make_memref_stream
        return f(nm, stype)

    def put_batch(self, s, a: np.ndarray):
        assert a.dtype == np.uint64 and a.ndim == 2 and a.flags["C_CONTIGUOUS"]
        p = a.ctypes.data
        self.lib.stream_emulator_put_memref_batch(s, p, p, 0, a.shape[0], a.shape[1], a.shape[1], 1, 0)

    def put(self, s, a: np.ndarray):
        assert a.dtype == np.uint64 and a.ndim == 1 and a.flags["C_CONTIGUOUS"]
        p = a.ctypes.data
        self.lib.stream_emulator_put_memref(s, p, p, 0, a.shape[0], 1, 0)

    def get_batch(self, s, out: np.ndarray):
        assert out.dtype == np.uint64 and out.ndim == 2 and out.flags["C_CONTIGUOUS"]
        p = out.ctypes.data
        self.lib.stream_emulator_get_memref_batch(s, p, p, 0, out.shape[0], out.shape[1], out.shape[1], 1)
        return out

    # ---- processes (batched forms)
    def keyswitch(self, sin, sout, p, ksk_index=0):
        self.lib.stream_emulator_make_memref_batched_keyswitch_lwe_u64_process(
            self.dfg, sin, sout, p.l_ks, p.b_ks, p.big_n, p.n, p.n + 1, ksk_index, self.rt.ctx)

    def bootstrap(self, sin, stlu, sout, p, bsk_index=0, mapped=False):
        f = (self.lib.stream_emulator_make_memref_batched_mapped_bootstrap_lwe_u64_process if mapped
             else self.lib.stream_emulator_make_memref_batched_bootstrap_lwe_u64_process)
        f(self.dfg, sin, stlu, sout, p.n, p.N, p.l, p.b, p.k, p.big_n + 1, bsk_index, self.rt.ctx)

    def add(self, a, b, out):
        self.lib.stream_emulator_make_memref_batched_add_lwe_ciphertexts_u64_process(self.dfg, a, b, out)

    def add_plaintext(self, a, b, out, cst=False):
        f = (self.lib.stream_emulator_make_memref_batched_add_plaintext_cst_lwe_ciphertext_u64_process
             if cst else self.lib.stream_emulator_make_memref_batched_add_plaintext_lwe_ciphertext_u64_process)
        f(self.dfg, a, b, out)

    def mul_cleartext(self, a, b, out, cst=False):
        f = (self.lib.stream_emulator_make_memref_batched_mul_cleartext_cst_lwe_ciphertext_u64_process
             if cst else self.lib.stream_emulator_make_memref_batched_mul_cleartext_lwe_ciphertext_u64_process)
        f(self.dfg, a, b, out)

    def negate(self, a, out):
        self.lib.stream_emulator_make_memref_batched_negate_lwe_ciphertext_u64_process(self.dfg, a, out)
