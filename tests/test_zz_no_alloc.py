"""GPU test of the allocation-free keyswitch entry point
(concrete_b200_keyswitch_lwe_ciphertext_vector_64_no_alloc, include/concrete_b200.h; csrc/keyswitch.cu).

No counterpart in the reference: the entry point exists for callers that may not allocate (CUDA
graph capture, serving loops).  It runs the tcgen05 INT8 keyswitch with its operand images in a
fixed `static __device__` array (72 MiB) instead of the stream-ordered pool.

Bar: every output word equal to cuda_keyswitch_lwe_ciphertext_vector_64 (the existing entry point,
bit-exact integer work, tolerance 0) and to the oracle, on calls in a row that share the array with
different shapes, so stale contents of an earlier call would show.  The file sorts last on purpose;
it passes alone in a fresh process (the new entry point is the
first thing the library runs here) and after the other tests."""
import ctypes as C

import numpy as np
import pytest

from harness import orc

pytestmark = pytest.mark.gpu

ENTRY = "concrete_b200_keyswitch_lwe_ciphertext_vector_64_no_alloc"
SCRATCH_BYTES = 72 << 20
BAD_ARGUMENT, UNSUPPORTED, TOO_LARGE = 1, 2, 3


def _case(seed, n_in, n_out, l_ks, num):
    rng = np.random.default_rng(seed)
    ksk = rng.integers(0, 2 ** 64, size=n_in * l_ks * (n_out + 1), dtype=np.uint64)
    cts = rng.integers(0, 2 ** 64, size=(num, n_in + 1), dtype=np.uint64)
    return ksk, cts


def _no_alloc(engine, d_cts, d_ksk, n_in, n_out, b_ks, l_ks, num, out=None):
    import torch

    if out is None:
        out = torch.full((num, n_out + 1), -1, dtype=torch.int64, device=engine.device)
    idx = engine.iota(max(num, 1))
    err = C.create_string_buffer(512)
    rc = getattr(engine.lib, ENTRY)(
        engine.stream, engine.gpu, out.data_ptr(), idx.data_ptr(), d_cts.data_ptr(), idx.data_ptr(),
        d_ksk.data_ptr(), n_in, n_out, b_ks, l_ks, num, err, len(err))
    return rc, err.value.decode(), out


def test_no_alloc_keyswitch_matches_existing_entry_point(engine):
    import torch
    from concrete_b200 import as_u64

    # ragged in all three GEMM dimensions; the second call is SMALLER in every dimension than the
    # first (its images cover a prefix of what the first left behind), the third is the
    # north-star keyswitch shape (P2048: 2048 x l 5 -> 742, key image 60 MiB of the 72) with more
    # rows than one digit chunk holds (2 chunks), the fourth repeats the second
    cases = [(11, 1024, 743, 4, 5, 300), (12, 600, 41, 3, 6, 77), (13, 2048, 742, 4, 5, 1300),
             (12, 600, 41, 3, 6, 77)]
    for i, (seed, n_in, n_out, b_ks, l_ks, num) in enumerate(cases):
        ksk, cts = _case(seed, n_in, n_out, l_ks, num)
        d_cts, d_ksk = engine.to_device(cts), engine.upload_ksk(ksk)
        rc, msg, out = _no_alloc(engine, d_cts, d_ksk, n_in, n_out, b_ks, l_ks, num)
        assert rc == 0, msg
        got = as_u64(out)
        existing = as_u64(engine.keyswitch(d_cts, d_ksk, n_in, n_out, b_ks, l_ks))
        assert np.array_equal(got, existing), f"call {i}: differs from the existing entry point"
        if num <= 300:
            assert np.array_equal(got, orc.keyswitch(cts, ksk, n_in, n_out, b_ks, l_ks))

    # the path holds no memory of its own: free device memory is the same before and after a call
    # (everything is loaded and warm by now; torch allocates `out` before the first reading)
    seed, n_in, n_out, b_ks, l_ks, num = cases[0]
    ksk, cts = _case(seed, n_in, n_out, l_ks, num)
    d_cts, d_ksk = engine.to_device(cts), engine.upload_ksk(ksk)
    out = torch.empty((num, n_out + 1), dtype=torch.int64, device=engine.device)
    torch.cuda.synchronize()
    free_before, _ = torch.cuda.mem_get_info(engine.gpu)
    rc, msg, out = _no_alloc(engine, d_cts, d_ksk, n_in, n_out, b_ks, l_ks, num, out=out)
    torch.cuda.synchronize()
    free_after, _ = torch.cuda.mem_get_info(engine.gpu)
    assert rc == 0, msg
    assert free_after == free_before
    assert np.array_equal(as_u64(out), orc.keyswitch(cts, ksk, n_in, n_out, b_ks, l_ks))


def test_no_alloc_keyswitch_rejects_what_does_not_fit(engine):
    import torch

    d = torch.zeros(16, dtype=torch.int64, device=engine.device)
    # key image alone: roundup(8 * 4097, 256) x roundup(4096 * 5, 128) = 640 MiB > 72 MiB
    rc, msg, _ = _no_alloc(engine, d, d, 4096, 4096, 4, 5, 1, out=d)
    assert rc == TOO_LARGE and str(SCRATCH_BYTES) in msg and "scratch" in msg
    # largest square-ish shape under the limit is accepted by the size check: P2048's
    Kp, Np = 2048 * 5, (743 * 8 + 255) // 256 * 256
    assert Np * Kp + 128 * Kp <= SCRATCH_BYTES
    # outside the exact INT8 range of the kernel (base_log 8; accumulator could wrap)
    rc, msg, _ = _no_alloc(engine, d, d, 64, 8, 8, 2, 1, out=d)
    assert rc == UNSUPPORTED and "base_log" in msg
    rc, msg, _ = _no_alloc(engine, d, d, 40000, 8, 7, 8, 1, out=d)
    assert rc == UNSUPPORTED
    # null pointer, and an empty batch is a no-op
    err = C.create_string_buffer(128)
    rc = getattr(engine.lib, ENTRY)(engine.stream, engine.gpu, None, d.data_ptr(), d.data_ptr(),
                                    d.data_ptr(), d.data_ptr(), 64, 8, 4, 2, 1, err, len(err))
    assert rc == BAD_ARGUMENT and b"null" in err.value
    rc, msg, _ = _no_alloc(engine, d, d, 64, 8, 4, 2, 0, out=d)
    assert rc == 0 and msg == ""
    torch.cuda.synchronize()
