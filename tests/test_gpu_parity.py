"""GPU parity tests: every call goes through the C ABI of libconcrete_b200.so and is compared with
the CPU oracle on the same seeded inputs:
  * keyswitch: every output word bit-exact;
  * one external product (one CMUX from an identical accumulator): every output WORD within the
    f64-FFT torus tolerance 8*sigma_fft;
  * full PBS: decryptions exact and every PHASE (b - <a,s>) within 8*sigma_fft of the oracle's.
    Word-level comparison of a full PBS is meaningless for any two f64 implementations — see
    harness.phase_distance.

Mirrors the reference's own GPU tests:
  TCB/tests_and_benchmarks/tests/test_keyswitch.cpp:106-171   (decrypt after keyswitch)
  TCB/tests_and_benchmarks/tests/test_classical_pbs.cpp:106-227 (decrypt after PBS, amortized +
                                                                 default entry points)
  TCB/tests_and_benchmarks/tests/test_fft.cpp:76-107          (FFT product vs schoolbook, 1e-9)
but with the stronger word-level pins.
"""
import ctypes as C

import numpy as np
import pytest

from harness import (KeySet, PARAMS, TfheParams, expand_lut, orc, phase_distance, phase_tolerance, sigma_fft_ulp,
                     torus_distance, trivial_glwe)

pytestmark = pytest.mark.gpu

_KEYS = {}


def keyset(name, n_override=None):
    key = (name, n_override)
    if key not in _KEYS:
        _KEYS[key] = KeySet(PARAMS[name], seed=42, n_override=n_override)
    return _KEYS[key]


def square_lut(p, bits):
    table = (np.arange(1 << bits) ** 2 % (1 << bits)).astype(np.uint64)
    return table, trivial_glwe(expand_lut(table, p.N, bits), p.k)


# ------------------------------------------------------------------------------------ keyswitch
@pytest.mark.parametrize("name,num", [("TOY512", 1), ("TOY512", 37), ("TOY256K2", 129),
                                      ("TOY1024", 300), ("P2048", 257), ("P4", 64), ("P6", 33)])
def test_keyswitch_bit_exact(engine, name, num):
    from concrete_b200 import as_u64

    ks = keyset(name)
    p = ks.p
    bits = min(p.msg_bits, 4)
    msgs = (np.arange(num) % (1 << bits)).astype(np.uint64)
    cts = ks.encrypt_big(msgs, bits, seed=11)
    ref = orc.keyswitch(cts, ks.ksk, p.big_n, p.n, p.b_ks, p.l_ks)
    got = as_u64(engine.keyswitch(engine.to_device(cts), engine.upload_ksk(ks.ksk), p.big_n, p.n,
                                  p.b_ks, p.l_ks))
    assert np.array_equal(got, ref)
    assert np.array_equal(ks.decrypt_small(got, bits), msgs)


@pytest.mark.parametrize("num", [64, 640])  # IMAD kernel / INT8 tensor-core kernel
def test_keyswitch_edge_values_and_indexes(engine, num):
    """a_i in {0, 2^63, 2^64-1, exact half-way rounding points}; permuted in/out index arrays."""
    import torch
    from concrete_b200 import as_u64

    ks = keyset("TOY512")
    p = ks.p
    rng = np.random.default_rng(5)
    cts = rng.integers(0, 2 ** 64, size=(num, p.big_n + 1), dtype=np.uint64)
    shift = 64 - p.b_ks * p.l_ks
    half = np.uint64(1) << np.uint64(shift - 1)
    cts[0, :] = 0
    cts[1, :] = np.uint64(1) << np.uint64(63)
    cts[2, :] = np.uint64(2 ** 64 - 1)
    cts[3, :] = half                       # exactly half-way: rounds up
    cts[4, :] = half - np.uint64(1)        # just below: rounds down
    cts[5, :] = (np.uint64(1) << np.uint64(63)) - half  # carry chain through every digit
    cts[6, ::2] = np.uint64(2 ** 64 - 1) - half + np.uint64(1)
    in_idx = rng.permutation(num).astype(np.uint64)
    out_idx = rng.permutation(num).astype(np.uint64)
    ref = np.empty((num, p.n + 1), dtype=np.uint64)
    ref[out_idx] = orc.keyswitch(np.ascontiguousarray(cts[in_idx]), ks.ksk, p.big_n, p.n, p.b_ks,
                                 p.l_ks)
    d_in_idx = torch.from_numpy(in_idx.view(np.int64)).to(engine.device)
    d_out_idx = torch.from_numpy(out_idx.view(np.int64)).to(engine.device)
    got = as_u64(engine.keyswitch(engine.to_device(cts), engine.upload_ksk(ks.ksk), p.big_n, p.n,
                                  p.b_ks, p.l_ks, in_idx=d_in_idx, out_idx=d_out_idx))
    assert np.array_equal(got, ref)


@pytest.mark.parametrize("n_in,n_out,b_ks,l_ks,num", [
    (8192, 915, 3, 6, 130),    # P7's shape: 48 kappa rows per tile (48 KiB + static shared memory)
    (16384, 1006, 4, 5, 40),   # P8's shape
    (700, 333, 1, 1, 5), (700, 333, 2, 12, 200), (513, 64, 2, 16, 129), (300, 63, 1, 48, 3),
    (1024, 599, 31, 2, 77), (2048, 742, 7, 7, 1), (96, 1000, 9, 7, 260),
    # batches >= 512 take the INT8 tensor-core path when digits fit s8 and s32 sums cannot wrap
    (2048, 742, 3, 5, 700), (2048, 801, 4, 3, 513), (16384, 1006, 4, 5, 520), (8192, 915, 3, 6, 640),
    (333, 15, 7, 9, 1025), (77, 130, 1, 3, 512), (1000, 500, 2, 20, 600),
    (700, 333, 9, 7, 600),      # digits do not fit s8 -> IMAD kernel at a large batch
    (40000, 16, 7, 9, 512)])    # (B/2) * 255 * K >= 2^31 -> IMAD kernel
def test_keyswitch_decomposition_shapes(engine, n_in, n_out, b_ks, l_ks, num):
    """Every tile configuration of both keyswitch kernels (IMAD: inputs per tile = min(8, 48 /
    l_ks), split-K on and off; INT8-MMA: ragged edges in all three GEMM dimensions).  The
    keyswitch is linear algebra over wrapping u64, so a uniformly random "KSK" checks the
    arithmetic word for word against the oracle."""
    from concrete_b200 import as_u64

    rng = np.random.default_rng(n_in + 31 * l_ks)
    ksk = rng.integers(0, 2 ** 64, size=n_in * l_ks * (n_out + 1), dtype=np.uint64)
    cts = rng.integers(0, 2 ** 64, size=(num, n_in + 1), dtype=np.uint64)
    ref = orc.keyswitch(cts, ksk, n_in, n_out, b_ks, l_ks)
    got = as_u64(engine.keyswitch(engine.to_device(cts), engine.upload_ksk(ksk), n_in, n_out,
                                  b_ks, l_ks))
    assert np.array_equal(got, ref)


@pytest.mark.parametrize("plan", [2])        # 2: the tcgen05 byte-plane GEMM whenever it is exact
@pytest.mark.parametrize("n_in,n_out,b_ks,l_ks,num", [
    (2048, 742, 3, 5, 700), (2048, 801, 4, 3, 129), (333, 15, 7, 9, 1025), (77, 130, 1, 3, 512),
    (1000, 500, 2, 20, 300), (16, 3, 4, 2, 5), (4096, 880, 4, 4, 257), (96, 17, 4, 3, 70001)])
def test_keyswitch_tensor_core_kernels(engine, plan, n_in, n_out, b_ks, l_ks, num):
    """The tensor-core keyswitch kernel forced (also below its automatic batch threshold):
    ragged edges in all three GEMM dimensions (samples not a multiple of 128, key columns not a
    multiple of the 256-column tile, K not a multiple of the 128-byte stage), more tiles than
    SMs (persistent loop, both accumulator buffers), one tile, > 65 535 rows.  Word for word
    against the oracle; random "KSK" (the keyswitch is linear algebra over wrapping u64)."""
    from concrete_b200 import as_u64

    rng = np.random.default_rng(n_in + 31 * l_ks + plan)
    ksk = rng.integers(0, 2 ** 64, size=n_in * l_ks * (n_out + 1), dtype=np.uint64)
    cts = rng.integers(0, 2 ** 64, size=(num, n_in + 1), dtype=np.uint64)
    ref = orc.keyswitch(cts, ksk, n_in, n_out, b_ks, l_ks)
    engine.lib.concrete_b200_set_ks_plan(plan)
    try:
        got = as_u64(engine.keyswitch(engine.to_device(cts), engine.upload_ksk(ksk), n_in, n_out,
                                      b_ks, l_ks))
    finally:
        engine.lib.concrete_b200_set_ks_plan(0)
    assert np.array_equal(got, ref)


def test_keyswitch_more_than_65535_rows(engine):
    """Small K (n_in * l_ks < 4096) and a batch above the 65 535 grid.y limit in ONE chunk of the
    INT8 tensor-core path: the digit kernel's row index must not sit on blockIdx.y."""
    from concrete_b200 import as_u64

    n_in, n_out, b_ks, l_ks, num = 96, 17, 4, 3, 70001
    rng = np.random.default_rng(99)
    ksk = rng.integers(0, 2 ** 64, size=n_in * l_ks * (n_out + 1), dtype=np.uint64)
    cts = rng.integers(0, 2 ** 64, size=(num, n_in + 1), dtype=np.uint64)
    ref = orc.keyswitch(cts, ksk, n_in, n_out, b_ks, l_ks)
    got = as_u64(engine.keyswitch(engine.to_device(cts), engine.upload_ksk(ksk), n_in, n_out,
                                  b_ks, l_ks))
    assert np.array_equal(got, ref)


def test_keyswitch_empty_batch(engine):
    import torch

    ks = keyset("TOY512")
    p = ks.p
    out = torch.full((4, p.n + 1), 7, dtype=torch.int64, device=engine.device)
    cts = torch.zeros((4, p.big_n + 1), dtype=torch.int64, device=engine.device)
    engine.keyswitch(cts, engine.upload_ksk(ks.ksk), p.big_n, p.n, p.b_ks, p.l_ks, out=out,
                     num_samples=0)
    engine.synchronize()
    assert bool((out == 7).all())


def test_keyswitch_golden_fixture(engine):
    """Committed fixture (tests/golden/make_golden.py): seeded inputs -> oracle output words."""
    import os
    from concrete_b200 import as_u64

    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "ks_pbs_toy512.npz"))
    ks = keyset("TOY512")
    p = ks.p
    got = as_u64(engine.keyswitch(engine.to_device(g["cts"]), engine.upload_ksk(ks.ksk), p.big_n,
                                  p.n, p.b_ks, p.l_ks))
    assert np.array_equal(got, g["ks_out"])


# ------------------------------------------------------------------------------------------ FFT
@pytest.mark.parametrize("N", [256, 512, 1024, 2048, 4096, 8192, 16384])
def test_fourier_polynomial_mul(engine, N):
    """test_fft.cpp:76-107: fwd x fwd -> pointwise -> inv equals the schoolbook negacyclic
    product within 1e-9 (doubles uniform in [-1, 1])."""
    import torch

    total = 6
    g = torch.Generator(device="cpu").manual_seed(N)
    a = torch.rand((total, N), dtype=torch.float64, generator=g) * 2 - 1
    b = torch.rand((total, N), dtype=torch.float64, generator=g) * 2 - 1
    got = engine.fourier_polynomial_mul(a.to(engine.device), b.to(engine.device)).cpu().numpy()
    an, bn = a.numpy(), b.numpy()
    for q in range(total):
        full = np.convolve(an[q], bn[q])
        ref = full[:N].copy()
        ref[: N - 1] -= full[N:]
        assert np.abs(got[q] - ref).max() < 1e-9


# ------------------------------------------------------------------------------------------ PBS
def _pbs_check(engine, ks, num, bits, lut_count=1, lut_stride=1, seed=13, use_ks=True):
    from concrete_b200 import as_u64

    p = ks.p
    table, lut = square_lut(p, bits)
    rng = np.random.default_rng(seed)
    msgs = rng.integers(0, 1 << bits, num).astype(np.uint64)
    if use_ks:
        mid = orc.keyswitch(ks.encrypt_big(msgs, bits, seed=seed), ks.ksk, p.big_n, p.n, p.b_ks,
                            p.l_ks)
    else:
        mid = ks.encrypt_small(msgs, bits, seed=seed)
    ref = orc.bootstrap(mid, lut, None, ks.fbsk, p.n, p.k, p.N, p.b, p.l)
    d_fbsk = engine.convert_bsk(ks.bsk, p.n, p.k, p.l, p.N)
    got = as_u64(engine.bootstrap(engine.to_device(mid), engine.to_device(lut), None, d_fbsk, p.n,
                                  p.k, p.N, p.b, p.l))
    assert np.array_equal(ks.decrypt_big(ref, bits), table[msgs]), "oracle itself is wrong"
    assert np.array_equal(ks.decrypt_big(got, bits), table[msgs])
    sig = sigma_fft_ulp(p)
    dist = phase_distance(ks, got, ref)
    assert dist.max() <= phase_tolerance(p), (np.log2(dist.max() + 1), np.log2(sig))
    # output noise of the GPU path is the oracle's, not larger (same keys, same inputs)
    expect = (table[msgs] << np.uint64(63 - bits)).astype(np.uint64)
    noise_got = (orc.lwe_decrypt(ks.sk_glwe, got) - expect).view(np.int64).astype(np.float64)
    noise_ref = (orc.lwe_decrypt(ks.sk_glwe, ref) - expect).view(np.int64).astype(np.float64)
    assert np.abs(noise_got).max() <= 2.0 * np.abs(noise_ref).max() + phase_tolerance(p)
    return got, ref


@pytest.mark.parametrize("name", ["TOY512", "TOY256K2", "TOY1024", "P2048", "P4", "P6", "P7",
                                  "P8"])
def test_external_product_word_tolerance(engine, name):
    """One CMUX from an identical, fully random GLWE accumulator (passed as the "LUT"; n = 1):
    the rotation, rounding and digits are exact integer work, so GPU and oracle feed identical
    digits to their FFTs and every output WORD must agree within the f64-FFT error budget
    8*sigma_fft(n=1).  This is the ciphertext-word tolerance the north star asks for."""
    import torch
    from concrete_b200 import as_u64

    ks = keyset(name, n_override=1 if name.startswith("P") else None)
    p = ks.p
    num = 12
    rng = np.random.default_rng(21)
    accs = rng.integers(0, 2 ** 64, size=(num, (p.k + 1) * p.N), dtype=np.uint64)
    cts = rng.integers(0, 2 ** 64, size=(num, 2), dtype=np.uint64)  # n = 1: (a_0, b)
    cts[0, 0] = 0                                     # skipped CMUX: pure rotation + extraction
    cts[1, 0] = np.uint64(1) << np.uint64(63)         # X^N = -1
    lut_idx = np.arange(num, dtype=np.uint64)
    one = TfheParams(**{**p.__dict__, "n": 1})
    fbsk1 = ks.fbsk[: (p.k + 1) ** 2 * p.l * (p.N // 2)]
    bsk1 = ks.bsk[: (p.k + 1) ** 2 * p.l * p.N]
    ref = orc.bootstrap(cts, accs, lut_idx, fbsk1, 1, p.k, p.N, p.b, p.l)
    dev = lambda a: torch.from_numpy(a.view(np.int64)).to(engine.device)  # noqa: E731
    d_fbsk = engine.convert_bsk(bsk1, 1, p.k, p.l, p.N)
    got = as_u64(engine.bootstrap(dev(cts), dev(accs), dev(lut_idx), d_fbsk, 1, p.k, p.N, p.b,
                                  p.l))
    dist = torus_distance(got, ref)
    sig = sigma_fft_ulp(one)
    assert dist.max() <= 8 * sig, (np.log2(dist.max() + 1), np.log2(sig))
    assert np.array_equal(got[0], ref[0])  # no CMUX executed: integer-only path, bit-exact
    signed = (got[1:] - ref[1:]).view(np.int64).astype(np.float64)
    assert abs(signed.mean()) <= sig / 16  # no systematic rounding bias (cf. functions.cuh:182)


@pytest.mark.parametrize("name,num,bits", [("TOY512", 37, 3), ("TOY256K2", 19, 2),
                                           ("TOY1024", 64, 3), ("P2048", 24, 4), ("P4", 16, 4)])
def test_pbs_parity(engine, name, num, bits):
    _pbs_check(engine, keyset(name), num, bits)


@pytest.mark.parametrize("name,n_small,bits", [("P6", 48, 4), ("P7", 32, 4), ("P8", 24, 4)])
def test_pbs_large_polynomials(engine, name, n_small, bits):
    """cfg 3 shapes (N = 4096 / 8192 / 16384) with a shortened small key so the oracle finishes in
    seconds; the n-iteration loop body is identical."""
    _pbs_check(engine, keyset(name, n_override=n_small), 8, bits, use_ks=False)


@pytest.mark.parametrize("k,N,b,l,sig_glwe", [
    (3, 512, 18, 1, 1.1919984450689246e-23 ** 0.5),   # BOOLEAN_DEFAULT / MESSAGE_1_CARRY_1
    (5, 256, 15, 1, 1.3996292326131784e-19 ** 0.5),   # MESSAGE_1_CARRY_0
    (2, 512, 8, 2, 1.6434266310406663e-15 ** 0.5),    # MESSAGE_2_CARRY_0
    (2, 1024, 23, 1, 8.645717832544903e-32 ** 0.5),   # BOOLEAN_TFHE_LIB / MESSAGE_*_CARRY_* rows
    (1, 4096, 22, 1, 4.70197740328915e-38 ** 0.5),    # MESSAGE_5_CARRY_0
    (1, 8192, 15, 2, 4.70197740328915e-38 ** 0.5),    # MESSAGE_3_CARRY_3
    (4, 256, 12, 2, 2.0 ** -40), (6, 256, 10, 3, 2.0 ** -40), (3, 2048, 20, 1, 2.0 ** -60),
    (2, 512, 17, 2, 2.0 ** -55), (1, 1024, 12, 3, 2.0 ** -55),  # base_log * level > 31: 64-bit states
])
def test_pbs_reference_parameter_table(engine, k, N, b, l, sig_glwe):
    """The (k, N, base_log, level) rows of the reference's own GPU test table
    (TCB/tests_and_benchmarks/tests/test_classical_pbs.cpp:232-376) plus the remaining (k, N)
    instantiations of the generic kernel, with the small dimension cut to 28 so that the oracle
    finishes in seconds (the CMUX loop body does not depend on n)."""
    prm = TfheParams(f"REF_k{k}_N{N}", 28, k, N, l, b, 3, 4, 2, 2.0 ** -24, sig_glwe)
    _pbs_check(engine, KeySet(prm, seed=42), 9, 2, use_ks=False)


@pytest.mark.parametrize("name,n_small", [("TOY512", None), ("P4", 24), ("P7", 16), ("P8", 8)])
def test_pbs_cluster_variant(engine, name, n_small):
    """The thread-block-cluster variant of the group kernel (one CTA per GLWE polynomial, Fourier
    digits exchanged through distributed shared memory; pbs_grp.cuh) is the default only at
    N = 16384 but must stay correct everywhere: same checks as the default path."""
    engine.lib.concrete_b200_set_pbs_plan(3)
    try:
        _pbs_check(engine, keyset(name, n_override=n_small), 7, 3, use_ks=False)
    finally:
        engine.lib.concrete_b200_set_pbs_plan(0)


@pytest.mark.parametrize("plan", [1, 2, 4])
@pytest.mark.parametrize("name,n_small", [("TOY1024", None), ("P2048", 20), ("P4", 24), ("P6", 16),
                                          ("P7", 12)])
def test_pbs_every_kernel_family(engine, name, n_small, plan):
    """Plans 1 (first generic kernel), 2 (group kernel) and 4 (v4: TMEM-resident accumulator copy,
    TMA key ring) on the shapes where the automatic plan would pick another family."""
    engine.lib.concrete_b200_set_pbs_plan(plan)
    try:
        _pbs_check(engine, keyset(name, n_override=n_small), 11, 3, use_ks=False)
    finally:
        engine.lib.concrete_b200_set_pbs_plan(0)


@pytest.mark.parametrize("k,N,b,l", [(1, 8192, 15, 2), (1, 1024, 12, 3), (1, 2048, 9, 3),
                                     (2, 1024, 10, 2), (1, 4096, 11, 2), (1, 2048, 16, 2),
                                     (1, 16384, 15, 2), (1, 16384, 22, 1), (1, 16384, 11, 3),
                                     (1, 16384, 9, 3)])
def test_pbs_v4_multi_level(engine, k, N, b, l):
    """v4 with l >= 2: Fourier accumulators parked in tensor memory between levels; 32-bit
    (b*l <= 31) and 64-bit (b*l > 31) decomposition states."""
    prm = TfheParams(f"V4_k{k}_N{N}_l{l}", 20, k, N, l, b, 3, 4, 2, 2.0 ** -24, 2.0 ** -55)
    _pbs_check(engine, KeySet(prm, seed=42), 13, 2, use_ks=False)


@pytest.mark.parametrize("name,n_small,alt,plan,num", [("P8", 10, 2, 0, 80), ("P8", 10, 1, 0, 9),
                                                       ("P4", 16, 3, 4, 450)])
def test_pbs_v4_tuning_variants(engine, name, n_small, alt, plan, num):
    """The instantiations behind concrete_b200_set_v4_alt stay correct: N = 16384 with 16 points per
    thread (the default has 32; last exchange by shuffle) and with packed decomposition states,
    k = 2 / N = 1024 with 8 points per thread.  Hook (and plan) are set before the key is converted:
    these variants have their own Fourier key layout."""
    engine.lib.concrete_b200_set_v4_alt(alt)
    engine.lib.concrete_b200_set_pbs_plan(plan)
    try:
        _pbs_check(engine, keyset(name, n_override=n_small), num, 3, use_ks=False)
    finally:
        engine.lib.concrete_b200_set_v4_alt(0)
        engine.lib.concrete_b200_set_pbs_plan(0)


def test_pbs_v4_rounds_and_tail(engine):
    """More ciphertexts than one wave of the persistent v4 grid (full rounds + a ragged tail spread
    over all CTAs) with the small dimension cut so that the oracle finishes in seconds."""
    ks = keyset("TOY1024", n_override=6)  # k = 1, N = 1024: 8 ciphertexts per CTA
    _pbs_check(engine, ks, 148 * 8 + 148 * 3 + 17, 3, use_ks=False)


def test_pbs_v4_cluster_rounds_and_tail(engine):
    """N = 16384 (one polynomial per SM, the two CTAs of a cluster per ciphertext): more ciphertexts
    than one wave of clusters plus a ragged tail, P8's decomposition, small dimension cut to 3."""
    ks = keyset("P8", n_override=12)
    _pbs_check(engine, ks, 74 * 2 + 29, 4, use_ks=False)


@pytest.mark.parametrize("num", [296, 297, 741 + 148 * 5])
def test_pbs_k2_batch_size_picks_kernel(engine, num):
    """k = 2, N = 1024 (P4): ONE converted key, two kernels - up to 296 samples the group kernel
    reads the v4 key layout (pbs_grp.cuh FOLD = 1), from 297 on the v4 kernel (one warp per
    polynomial, 5 ciphertexts per CTA) runs.  Both sides of the switch against the oracle."""
    _pbs_check(engine, keyset("P4", n_override=8), num, 3, use_ks=False)


@pytest.mark.parametrize("num", [9, 760])
@pytest.mark.parametrize("b", [17, 10])
def test_pbs_k2_two_levels_both_kernels(engine, num, b):
    """Same switch with two levels: base_log * level > 31 (64-bit decomposition states) and <= 31
    (32-bit states, parked in tensor memory between the levels by the v4 kernel)."""
    prm = TfheParams(f"K2_N1024_b{b}", 6, 2, 1024, 2, b, 3, 4, 2, 2.0 ** -24, 2.0 ** -55)
    _pbs_check(engine, KeySet(prm, seed=42), num, 2, use_ks=False)


@pytest.mark.parametrize("name,n_small,num", [("P8", 12, 90), ("P7", 12, 160), ("P2048", 20, 900),
                                              ("P4", 16, 750)])
def test_pbs_deterministic_and_families_agree(engine, name, n_small, num):
    """Race detector + cross-check of independently written kernels: the default kernel run twice
    on the same inputs must give bit-identical outputs (no ordering hazard in the shared-memory /
    DSMEM / TMEM / key-ring hand-offs), and its phases must agree with the first generic kernel
    (plan 1: different thread mapping, different transform passes, no TMA, no TMEM) more tightly
    than either agrees with the CPU oracle (measured on B200: rms 2^39..2^41.5 between kernel
    families at these sizes, 2^42 against the oracle; tools/p8_phase_stats.py)."""
    from concrete_b200 import as_u64

    ks = keyset(name, n_override=n_small)
    p = ks.p
    bits = 3
    table, lut = square_lut(p, bits)
    rng = np.random.default_rng(99)
    msgs = rng.integers(0, 1 << bits, num).astype(np.uint64)
    d_mid = engine.to_device(ks.encrypt_small(msgs, bits, seed=99))
    d_lut = engine.to_device(lut)
    outs = []
    try:
        for plan in (0, 0, 1):
            engine.lib.concrete_b200_set_pbs_plan(plan)
            d_fbsk = engine.convert_bsk(ks.bsk, p.n, p.k, p.l, p.N)
            outs.append(as_u64(engine.bootstrap(d_mid, d_lut, None, d_fbsk, p.n, p.k, p.N, p.b, p.l)))
    finally:
        engine.lib.concrete_b200_set_pbs_plan(0)
    assert np.array_equal(outs[0], outs[1]), "default kernel is not deterministic"
    for o in (outs[0], outs[2]):
        assert np.array_equal(ks.decrypt_big(o, bits), table[msgs])
    dist = phase_distance(ks, outs[0], outs[2])
    assert dist.max() <= phase_tolerance(p) / 2, (np.log2(dist.max() + 1), np.log2(phase_tolerance(p)))


def test_pbs_empty_batch(engine):
    """num_samples == 0 is a no-op (no launch, output untouched), like the keyswitch."""
    import torch

    ks = keyset("TOY512")
    p = ks.p
    out = torch.full((3, p.big_n + 1), 5, dtype=torch.int64, device=engine.device)
    cts = torch.zeros((3, p.n + 1), dtype=torch.int64, device=engine.device)
    lut = engine.to_device(trivial_glwe(np.zeros(p.N, dtype=np.uint64), p.k))
    d_fbsk = engine.convert_bsk(ks.bsk, p.n, p.k, p.l, p.N)
    engine.bootstrap(cts, lut, None, d_fbsk, p.n, p.k, p.N, p.b, p.l, out=out, num_samples=0)
    engine.synchronize()
    assert bool((out == 5).all())


def test_pbs_zero_mask_elements_and_indexes(engine):
    """Mask elements equal to 0 are skipped (bootstrap.rs:287); permuted index arrays; per-sample
    LUT indexes into a table of two accumulators."""
    import torch
    from concrete_b200 import as_u64

    ks = keyset("TOY512")
    p = ks.p
    bits, num = 3, 16
    t0 = (np.arange(8) ** 2 % 8).astype(np.uint64)
    t1 = ((np.arange(8) * 3 + 1) % 8).astype(np.uint64)
    luts = np.stack([trivial_glwe(expand_lut(t, p.N, bits), p.k) for t in (t0, t1)])
    msgs = (np.arange(num) % 8).astype(np.uint64)
    cts = ks.encrypt_small(msgs, bits)
    # zeroing mask elements whose key bit is 0 leaves the plaintext unchanged
    zero_pos = np.where(ks.sk_small == 0)[0][:5]
    cts[:, zero_pos] = 0
    lut_idx = (np.arange(num) % 2).astype(np.uint64)
    rng = np.random.default_rng(1)
    in_idx = rng.permutation(num).astype(np.uint64)
    out_idx = rng.permutation(num).astype(np.uint64)
    ref = np.empty((num, p.big_n + 1), dtype=np.uint64)
    ref[out_idx] = orc.bootstrap(np.ascontiguousarray(cts[in_idx]), luts, lut_idx, ks.fbsk, p.n,
                                 p.k, p.N, p.b, p.l)
    dev = lambda a: torch.from_numpy(a.view(np.int64)).to(engine.device)  # noqa: E731
    d_fbsk = engine.convert_bsk(ks.bsk, p.n, p.k, p.l, p.N)
    got = as_u64(engine.bootstrap(dev(cts), dev(luts), dev(lut_idx), d_fbsk, p.n, p.k, p.N, p.b,
                                  p.l, in_idx=dev(in_idx), out_idx=dev(out_idx)))
    assert phase_distance(ks, got, ref).max() <= phase_tolerance(p)
    expect = np.empty(num, dtype=np.uint64)
    sel = np.where(lut_idx == 0, t0[msgs[in_idx]], t1[msgs[in_idx]])
    expect[out_idx] = sel
    assert np.array_equal(ks.decrypt_big(got, bits), expect)


def test_pbs_many_lut(engine):
    """lut_count > 1: sample extraction at i*lut_stride into out + i*num_samples*(kN+1)
    (programmable_bootstrap_classic.cuh:211-259)."""
    from concrete_b200 import as_u64

    ks = keyset("TOY512")
    p = ks.p
    bits, num, lut_count, lut_stride = 3, 9, 3, 5
    table, lut = square_lut(p, bits)
    msgs = (np.arange(num) % 8).astype(np.uint64)
    cts = ks.encrypt_small(msgs, bits)
    d_fbsk = engine.convert_bsk(ks.bsk, p.n, p.k, p.l, p.N)
    got = as_u64(engine.bootstrap(engine.to_device(cts), engine.to_device(lut), None, d_fbsk, p.n,
                                  p.k, p.N, p.b, p.l, lut_count=lut_count, lut_stride=lut_stride))
    got = got.reshape(lut_count, num, p.big_n + 1)
    for i in range(lut_count):
        ref = orc.bootstrap(cts, lut, None, ks.fbsk, p.n, p.k, p.N, p.b, p.l, nth=i * lut_stride)
        assert phase_distance(ks, got[i], ref).max() <= phase_tolerance(p)


def test_pbs_golden_fixture(engine):
    import os
    from concrete_b200 import as_u64

    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "ks_pbs_toy512.npz"))
    ks = keyset("TOY512")
    p = ks.p
    d_fbsk = engine.convert_bsk(ks.bsk, p.n, p.k, p.l, p.N)
    got = as_u64(engine.bootstrap(engine.to_device(g["ks_out"]), engine.to_device(g["lut"]), None,
                                  d_fbsk, p.n, p.k, p.N, p.b, p.l))
    assert phase_distance(ks, got, g["pbs_out"]).max() <= phase_tolerance(p)
    assert np.array_equal(ks.decrypt_big(got, 3), g["expected"])


def test_bsk_convert_matches_oracle_up_to_layout(engine):
    """The Fourier layout is private, so compare through its only consumer: an external product
    with the converted key equals the oracle's (checked by the PBS tests), and converting twice
    gives identical bytes (determinism)."""
    import torch

    ks = keyset("TOY1024")
    p = ks.p
    a = engine.convert_bsk(ks.bsk, p.n, p.k, p.l, p.N)
    b = engine.convert_bsk(ks.bsk, p.n, p.k, p.l, p.N)
    assert torch.equal(a, b)
    assert a.numel() == ks.bsk.size  # fits the caller-allocated n(k+1)^2 l N * 8 bytes


# ---------------------------------------------------------------------- chain, legacy, linalg
def test_keyswitch_bootstrap_chain_matches_two_calls(engine):
    import torch

    ks = keyset("TOY1024")
    p = ks.p
    bits, num = 3, 50
    table, lut = square_lut(p, bits)
    msgs = (np.arange(num) % 8).astype(np.uint64)
    d_cts = engine.to_device(ks.encrypt_big(msgs, bits))
    d_ksk = engine.upload_ksk(ks.ksk)
    d_fbsk = engine.convert_bsk(ks.bsk, p.n, p.k, p.l, p.N)
    d_lut = engine.to_device(lut)
    mid = engine.keyswitch(d_cts, d_ksk, p.big_n, p.n, p.b_ks, p.l_ks)
    two = engine.bootstrap(mid, d_lut, None, d_fbsk, p.n, p.k, p.N, p.b, p.l)
    one = engine.keyswitch_bootstrap(d_cts, d_ksk, d_fbsk, d_lut, None, p)
    engine.synchronize()
    assert torch.equal(one, two)  # same kernels, same order: identical bits


def test_legacy_entry_points(engine):
    """cuda_bootstrap_{low_latency,amortized}_lwe_ciphertext_vector_64 and the 10-argument
    keyswitch (backends/concrete-cuda/rust_api/src/cuda_bind.rs:187-205,320-331)."""
    import torch
    from concrete_b200 import as_u64

    lib = engine.lib
    ks = keyset("TOY512")
    p = ks.p
    bits, num = 3, 11
    table, lut = square_lut(p, bits)
    msgs = (np.arange(num) % 8).astype(np.uint64)
    cts = ks.encrypt_big(msgs, bits)
    d_cts, d_ksk = engine.to_device(cts), engine.upload_ksk(ks.ksk)
    d_mid = torch.empty((num, p.n + 1), dtype=torch.int64, device=engine.device)
    lib.cuda_keyswitch_lwe_ciphertext_vector_64_legacy(
        engine.stream, 0, d_mid.data_ptr(), d_cts.data_ptr(), d_ksk.data_ptr(), p.big_n, p.n,
        p.b_ks, p.l_ks, num)
    ref_mid = orc.keyswitch(cts, ks.ksk, p.big_n, p.n, p.b_ks, p.l_ks)
    assert np.array_equal(as_u64(d_mid), ref_mid)

    d_fbsk = torch.empty(ks.bsk.size, dtype=torch.int64, device=engine.device)
    lib.cuda_convert_lwe_bootstrap_key_64(d_fbsk.data_ptr(), ks.bsk.ctypes.data, engine.stream, 0,
                                          p.n, p.k, p.l, p.N)
    lib.cuda_synchronize_stream(engine.stream, 0)
    d_lut = engine.to_device(lut)
    # lwe_idx offsets into lut_vector_indexes (cuda_bind.rs:187-205)
    d_lut_idx = torch.zeros(num + 4, dtype=torch.int64, device=engine.device)
    ref = orc.bootstrap(ref_mid, lut, None, ks.fbsk, p.n, p.k, p.N, p.b, p.l)
    for scratch, run, cleanup, has_level in (
            (lib.scratch_cuda_bootstrap_low_latency_64,
             lib.cuda_bootstrap_low_latency_lwe_ciphertext_vector_64,
             lib.cleanup_cuda_bootstrap_low_latency, True),
            (lib.scratch_cuda_bootstrap_amortized_64,
             lib.cuda_bootstrap_amortized_lwe_ciphertext_vector_64,
             lib.cleanup_cuda_bootstrap_amortized, False)):
        buf = C.POINTER(C.c_int8)()
        if has_level:
            scratch(engine.stream, 0, C.byref(buf), p.k, p.N, p.l, num, 49152, True)
        else:
            scratch(engine.stream, 0, C.byref(buf), p.k, p.N, num, 49152, True)
        d_out = torch.empty((num, p.big_n + 1), dtype=torch.int64, device=engine.device)
        run(engine.stream, 0, d_out.data_ptr(), d_lut.data_ptr(), d_lut_idx.data_ptr(),
            d_mid.data_ptr(), d_fbsk.data_ptr(), buf, p.n, p.k, p.N, p.b, p.l, num, 1, 4, 49152)
        lib.cuda_synchronize_stream(engine.stream, 0)
        got = as_u64(d_out)
        assert phase_distance(ks, got, ref).max() <= phase_tolerance(p)
        assert np.array_equal(ks.decrypt_big(got, bits), table[msgs])
        cleanup(engine.stream, 0, C.byref(buf))
        assert not buf


def test_amortized_current_names(engine):
    import torch
    from concrete_b200 import as_u64

    lib = engine.lib
    ks = keyset("TOY256K2")
    p = ks.p
    bits, num = 2, 7
    table, lut = square_lut(p, bits)
    msgs = (np.arange(num) % 4).astype(np.uint64)
    cts = ks.encrypt_small(msgs, bits)
    d_fbsk = engine.convert_bsk(ks.bsk, p.n, p.k, p.l, p.N)
    d_cts, d_lut = engine.to_device(cts), engine.to_device(lut)
    d_out = torch.empty((num, p.big_n + 1), dtype=torch.int64, device=engine.device)
    idx = engine.iota(num)
    zeros = torch.zeros(num, dtype=torch.int64, device=engine.device)
    buf = C.POINTER(C.c_int8)()
    lib.scratch_cuda_programmable_bootstrap_amortized_64(engine.stream, 0, C.byref(buf), p.k, p.N,
                                                         num, True)
    lib.cuda_programmable_bootstrap_amortized_lwe_ciphertext_vector_64(
        engine.stream, 0, d_out.data_ptr(), idx.data_ptr(), d_lut.data_ptr(), zeros.data_ptr(),
        d_cts.data_ptr(), idx.data_ptr(), d_fbsk.data_ptr(), buf, p.n, p.k, p.N, p.b, p.l, num)
    lib.cuda_synchronize_stream(engine.stream, 0)
    lib.cleanup_cuda_programmable_bootstrap_amortized(engine.stream, 0, C.byref(buf))
    got = as_u64(d_out)
    assert np.array_equal(ks.decrypt_big(got, bits), table[msgs])


def test_linear_algebra(engine):
    """TCB/include/linear_algebra.h:15-47 — wrapping u64 elementwise ops, bit-exact."""
    import torch
    from concrete_b200 import as_u64

    lib = engine.lib
    n, cnt = 37, 23
    rng = np.random.default_rng(2)
    a = rng.integers(0, 2 ** 64, size=(cnt, n + 1), dtype=np.uint64)
    b = rng.integers(0, 2 ** 64, size=(cnt, n + 1), dtype=np.uint64)
    pt = rng.integers(0, 2 ** 64, size=cnt, dtype=np.uint64)
    d_a, d_b, d_pt = engine.to_device(a), engine.to_device(b), engine.to_device(pt)
    out = torch.empty_like(d_a)
    lib.cuda_add_lwe_ciphertext_vector_64(engine.stream, 0, out.data_ptr(), d_a.data_ptr(),
                                          d_b.data_ptr(), n, cnt)
    assert np.array_equal(as_u64(out), a + b)
    lib.cuda_negate_lwe_ciphertext_vector_64(engine.stream, 0, out.data_ptr(), d_a.data_ptr(), n,
                                             cnt)
    assert np.array_equal(as_u64(out), np.uint64(0) - a)
    lib.cuda_add_lwe_ciphertext_vector_plaintext_vector_64(
        engine.stream, 0, out.data_ptr(), d_a.data_ptr(), d_pt.data_ptr(), n, cnt)
    ref = a.copy()
    ref[:, n] += pt
    assert np.array_equal(as_u64(out), ref)
    lib.cuda_mult_lwe_ciphertext_vector_cleartext_vector_64(
        engine.stream, 0, out.data_ptr(), d_a.data_ptr(), d_pt.data_ptr(), n, cnt)
    assert np.array_equal(as_u64(out), a * pt[:, None])


# ----------------------------------------------------------------------- runtime wrapper layer
def test_memref_wrappers_host_buffers(engine):
    """memref_batched_{keyswitch,bootstrap,mapped_bootstrap}_lwe[_cuda]_u64 with HOST buffers, as
    Concrete's generated code calls them (wrappers.cpp:110-363,745-826)."""
    from concrete_b200.runtime import Runtime

    ks = keyset("TOY1024")
    p = ks.p
    bits, num = 3, 40
    table = (np.arange(8) ** 2 % 8).astype(np.uint64)
    lut_poly = expand_lut(table, p.N, bits)
    msgs = (np.arange(num) % 8).astype(np.uint64)
    cts = ks.encrypt_big(msgs, bits)
    rt = Runtime(num_gpus=1)
    ki, bi = rt.add_ksk(ks.ksk), rt.add_bsk(ks.bsk)
    ref_mid = orc.keyswitch(cts, ks.ksk, p.big_n, p.n, p.b_ks, p.l_ks)
    ref_out = orc.bootstrap(ref_mid, trivial_glwe(lut_poly, p.k), None, ks.fbsk, p.n, p.k, p.N,
                            p.b, p.l)
    for cpu_named in (False, True):
        mid = np.zeros((num, p.n + 1), dtype=np.uint64)
        out = np.zeros((num, p.big_n + 1), dtype=np.uint64)
        rt.batched_keyswitch(mid, cts, p.l_ks, p.b_ks, p.big_n, p.n, ki, cpu_named=cpu_named)
        assert np.array_equal(mid, ref_mid)
        rt.batched_bootstrap(out, mid, lut_poly, p.n, p.N, p.l, p.b, p.k, bi, cpu_named=cpu_named)
        assert phase_distance(ks, out, ref_out).max() <= phase_tolerance(p)
        assert np.array_equal(ks.decrypt_big(out, bits), table[msgs])
    # mapped: one table per sample
    tables = np.stack([np.roll(table, s % 8) for s in range(num)])
    tlus = np.stack([expand_lut(t, p.N, bits) for t in tables])
    out = np.zeros((num, p.big_n + 1), dtype=np.uint64)
    rt.batched_mapped_bootstrap(out, ref_mid, tlus, p.n, p.N, p.l, p.b, p.k, bi)
    assert np.array_equal(ks.decrypt_big(out, bits), tables[np.arange(num), msgs])
    # single-ciphertext forms and the fused chain
    one_mid = np.zeros(p.n + 1, dtype=np.uint64)
    rt.keyswitch(one_mid, cts[3], p.l_ks, p.b_ks, p.big_n, p.n, ki)
    assert np.array_equal(one_mid, ref_mid[3])
    one_out = np.zeros(p.big_n + 1, dtype=np.uint64)
    rt.bootstrap(one_out, one_mid, lut_poly, p.n, p.N, p.l, p.b, p.k, bi)
    assert ks.decrypt_big(one_out[None, :], bits)[0] == table[msgs[3]]
    out2 = np.zeros((num, p.big_n + 1), dtype=np.uint64)
    rt.batched_keyswitch_bootstrap(out2, cts, lut_poly, p, ki, bi)
    assert np.array_equal(ks.decrypt_big(out2, bits), table[msgs])
    rt.close()


def test_memref_wrappers_pipelined_chunks(engine):
    """Batches large enough for the wrappers' copy/compute pipeline (several chunks, ragged last
    chunk; wrappers.cpp::run_pipelined): the keyswitch words equal the oracle's for every sample,
    every PBS decrypts to LUT[m], and the mapped form picks each sample's own table across chunk
    boundaries."""
    from concrete_b200.runtime import Runtime

    ks = keyset("TOY1024")
    p = ks.p
    bits, num = 3, 4 * 888 + 333
    table = (np.arange(8) ** 2 % 8).astype(np.uint64)
    lut_poly = expand_lut(table, p.N, bits)
    rng = np.random.default_rng(17)
    msgs = rng.integers(0, 8, num).astype(np.uint64)
    cts = ks.encrypt_big(msgs, bits)
    rt = Runtime(num_gpus=1)
    ki, bi = rt.add_ksk(ks.ksk), rt.add_bsk(ks.bsk)
    ref_mid = orc.keyswitch(cts, ks.ksk, p.big_n, p.n, p.b_ks, p.l_ks)
    mid = np.zeros((num, p.n + 1), dtype=np.uint64)
    out = np.zeros((num, p.big_n + 1), dtype=np.uint64)
    rt.batched_keyswitch(mid, cts, p.l_ks, p.b_ks, p.big_n, p.n, ki)
    assert np.array_equal(mid, ref_mid)
    rt.batched_bootstrap(out, mid, lut_poly, p.n, p.N, p.l, p.b, p.k, bi)
    assert np.array_equal(ks.decrypt_big(out, bits), table[msgs])
    tables = np.stack([np.roll(table, s % 8) for s in range(num)])
    tlus = np.stack([expand_lut(t, p.N, bits) for t in tables])
    out = np.zeros((num, p.big_n + 1), dtype=np.uint64)
    rt.batched_mapped_bootstrap(out, ref_mid, tlus, p.n, p.N, p.l, p.b, p.k, bi)
    assert np.array_equal(ks.decrypt_big(out, bits), tables[np.arange(num), msgs])
    rt.close()


# ----------------------------------------------------------- full-size properties (cfg 2 shape)
def test_full_size_batch_properties(engine):
    """BASELINE.json configs[1] size (16 384 ciphertexts, P2048): too big for the oracle, so check
    size-independent properties: every decryption equals LUT[m]; permuting the batch permutes the
    keyswitch output words identically; a spot sample of ciphertexts matches the oracle."""
    import torch
    from concrete_b200 import as_u64

    ks = keyset("P2048")
    p = ks.p
    bits, num = 4, 16384
    table, lut = square_lut(p, bits)
    rng = np.random.default_rng(99)
    msgs = rng.integers(0, 1 << bits, num).astype(np.uint64)
    cts = ks.encrypt_big(msgs, bits, seed=77)
    d_cts = engine.to_device(cts)
    d_ksk = engine.upload_ksk(ks.ksk)
    d_fbsk = engine.convert_bsk(ks.bsk, p.n, p.k, p.l, p.N)
    mid = engine.keyswitch(d_cts, d_ksk, p.big_n, p.n, p.b_ks, p.l_ks)
    out = engine.bootstrap(mid, engine.to_device(lut), None, d_fbsk, p.n, p.k, p.N, p.b, p.l)
    engine.synchronize()
    assert np.array_equal(ks.decrypt_big(as_u64(out), bits), table[msgs])
    # permutation equivariance of the keyswitch (bit-exact)
    perm = torch.from_numpy(rng.permutation(num)).to(engine.device)
    mid_perm = engine.keyswitch(d_cts[perm].contiguous(), d_ksk, p.big_n, p.n, p.b_ks, p.l_ks)
    assert torch.equal(mid_perm, mid[perm])
    # spot sample against the oracle
    pick = rng.choice(num, 16, replace=False)
    ref_mid = orc.keyswitch(np.ascontiguousarray(cts[pick]), ks.ksk, p.big_n, p.n, p.b_ks, p.l_ks)
    assert np.array_equal(as_u64(mid)[pick], ref_mid)
    ref_out = orc.bootstrap(ref_mid, lut, None, ks.fbsk, p.n, p.k, p.N, p.b, p.l)
    assert phase_distance(ks, as_u64(out)[pick], ref_out).max() <= phase_tolerance(p)
