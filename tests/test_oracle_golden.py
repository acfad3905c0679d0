"""Pins of the CPU oracle (oracle/tfhe_oracle.cpp) against the reference's own vectors and tests.

Golden vectors are the reference's doc-test values (paths under
/root/reference/third_party/tfhe-rs/tfhe/src/core_crypto):
  commons/math/decomposition/decomposer.rs:116-122   closest_representable(1_340_987_234u32)
  algorithms/polynomial_algorithms.rs:491-499         monomial div  [1,2,3] / X^2
  algorithms/polynomial_algorithms.rs:551-559         monomial mul  [1,2,3] * X^2
  algorithms/slice_algorithms.rs:355-362              slice_wrapping_sub_scalar_mul_assign
Property tests restate
  commons/math/decomposition/tests.rs:32-125          decompose/recompose, rounding idempotence
  fft_impl/fft64/math/fft/tests.rs:6-76,79-219         FFT roundtrip < 2^(64-50), product == schoolbook
  algorithms/polynomial_algorithms.rs:1265-1300        monomial mul/div roundtrip
  algorithms/test/lwe_keyswitch.rs:8-100               decrypt after keyswitch
  algorithms/test/lwe_programmable_bootstrapping.rs:69-155  decrypt after PBS == f(msg)
and a pure-Python restatement of the keyswitch cross-checks the C++ one word for word.
"""
import os

import numpy as np
import pytest

import oracle as orc
from harness import KeySet, PARAMS, decode, encode, expand_lut, trivial_glwe

M64 = (1 << 64) - 1


# ------------------------------------------------------------------ reference doc-test vectors
def test_closest_representable_doc_vector():
    assert orc.closest_representable(1_340_987_234, 3, 4, bits=32) == 1_341_128_704


def test_monomial_doc_vectors():
    poly = np.array([1, 2, 3], dtype=np.uint8)
    assert orc.monomial_div(poly, 2).tolist() == [3, 255, 254]
    assert orc.monomial_mul(poly, 2).tolist() == [254, 253, 1]


def test_sub_scalar_mul_doc_vector():
    # the keyswitch inner loop primitive; restated with numpy wrapping arithmetic and checked
    # through a 1-level, 1-input keyswitch whose single digit is `scalar`
    first = np.array([1, 2, 3, 4, 5, 6], dtype=np.uint8)
    second = np.array([255, 255, 255, 1, 2, 3], dtype=np.uint8)
    assert (first - second * np.uint8(4)).astype(np.uint8).tolist() == [5, 6, 7, 0, 253, 250]
    # u64 version through orc.keyswitch: n_in=1, level=1, base_log=4, a_0 = 4<<60 -> digit 4
    ksk = np.array([2 ** 64 - 1, 2 ** 64 - 1, 2 ** 64 - 1, 1, 2, 3], dtype=np.uint64)  # n_out=5
    ct = np.array([[4 << 60, 6]], dtype=np.uint64)
    out = orc.keyswitch(ct, ksk, 1, 5, 4, 1)
    expect = (np.array([0, 0, 0, 0, 0, 6], dtype=np.uint64) - ksk * np.uint64(4))
    assert np.array_equal(out[0], expect)


# ---------------------------------------------------------------------- decomposition properties
def _valid_decomposers(bits):
    return [(b, l) for b in range(1, bits) for l in range(1, bits) if b * l < bits and b <= 31]


@pytest.mark.parametrize("bits", [32, 64])
def test_decompose_recompose(bits):
    rng = np.random.default_rng(bits)
    mod = 1 << bits
    for (b, l) in _valid_decomposers(bits)[::7]:
        for _ in range(6):
            x = int(rng.integers(0, mod, dtype=np.uint64 if bits == 64 else np.uint32))
            terms = orc.decompose(x, l, b, bits)  # index 0 = level l
            half = 1 << (b - 1)
            rec = 0
            for idx, t in enumerate(terms):
                level = l - idx
                assert -half <= t <= half
                rec += t << (bits - b * level)
            assert rec % mod == orc.closest_representable(x, l, b, bits)


@pytest.mark.parametrize("bits", [32, 64])
def test_round_to_closest_representable(bits):
    rng = np.random.default_rng(bits + 1)
    mod = 1 << bits
    for (b, l) in _valid_decomposers(bits)[::5]:
        if bits - b * l - 1 < 1:
            continue
        x = int(rng.integers(0, mod, dtype=np.uint64 if bits == 64 else np.uint32))
        r = orc.closest_representable(x, l, b, bits)
        eps = (1 << (bits - b * l - 1)) // 2
        assert orc.closest_representable((r + eps) % mod, l, b, bits) == r
        assert orc.closest_representable((r - eps) % mod, l, b, bits) == r
        assert orc.closest_representable(r, l, b, bits) == r


def test_modulus_switch_and_from_torus():
    # fft_impl/common.rs:17-25: round to log_modulus bits
    assert orc.modulus_switch(0, 12) == 0
    assert orc.modulus_switch(1 << 52, 12) == 1
    assert orc.modulus_switch((1 << 51), 12) == 1          # half rounds up
    assert orc.modulus_switch((1 << 51) - 1, 12) == 0
    assert orc.modulus_switch(M64, 12) == 0                 # wrapping_add: rounds up to 2N == 0
    # commons/math/torus/mod.rs:72-78
    assert orc.from_torus(0.25) == 1 << 62
    assert orc.from_torus(-0.25) == (M64 + 1) - (1 << 62)
    assert orc.from_torus(3.0) == 0
    assert orc.from_torus(1.5) in ((1 << 63), (1 << 63) - 1, M64 - (1 << 63) + 1)


# --------------------------------------------------------------------------- polynomial helpers
def test_monomial_mul_div_roundtrip():
    rng = np.random.default_rng(0)
    for N in (8, 64, 1024):
        poly = rng.integers(0, 2 ** 64, size=N, dtype=np.uint64)
        for r in (int(rng.integers(0, N)), int(rng.integers(0, N)) + 2048, 0, N, 2 * N - 1):
            assert np.array_equal(orc.monomial_div(orc.monomial_mul(poly, r), r), poly)
            assert np.array_equal(orc.monomial_mul(orc.monomial_div(poly, r), r), poly)
            assert np.array_equal(orc.monomial_mul_and_subtract(poly, r),
                                  orc.monomial_mul(poly, r) - poly)


def test_sample_extract_is_constant_coefficient_of_decryption():
    # glwe_sample_extraction.rs:90-165: LWE(nth) decrypts to coefficient nth of the GLWE plaintext
    rng = np.random.default_rng(3)
    k, N = 2, 64
    key = rng.integers(0, 2, size=k * N, dtype=np.uint64)
    mask = rng.integers(0, 2 ** 64, size=k * N, dtype=np.uint64)
    msg = rng.integers(0, 2 ** 64, size=N, dtype=np.uint64)
    body = msg.copy()
    for p in range(k):
        body += orc.negacyclic_mul_schoolbook(mask[p * N:(p + 1) * N], key[p * N:(p + 1) * N])
    glwe = np.concatenate([mask, body])
    for nth in (0, 1, 17, N - 1):
        lwe = orc.sample_extract(glwe, k, N, nth)
        dec = orc.lwe_decrypt(key, lwe[None, :])[0]
        assert dec == msg[nth]


def test_negacyclic_mul_binary_matches_schoolbook():
    rng = np.random.default_rng(4)
    for N in (64, 512):
        a = rng.integers(0, 2 ** 64, size=N, dtype=np.uint64)
        s = rng.integers(0, 2, size=N, dtype=np.uint64)
        assert np.array_equal(orc.negacyclic_mul_binary(a, s), orc.negacyclic_mul_schoolbook(a, s))


# ------------------------------------------------------------------------------------------ FFT
@pytest.mark.parametrize("size_log", list(range(5, 15)))
def test_fft_roundtrip(size_log):
    N = 1 << size_log
    rng = np.random.default_rng(size_log)
    poly = rng.integers(0, 2 ** 64, size=N, dtype=np.uint64)
    f = orc.fft_forward_as_torus(poly)
    back = np.zeros(N, dtype=np.uint64)
    orc.fft_add_backward_as_torus(back, f)
    d = (poly - back).view(np.int64)
    assert np.abs(d).max() < (1 << (64 - 50))


@pytest.mark.parametrize("size_log", [5, 8, 10])
def test_fft_product_vs_schoolbook(size_log):
    # fft/tests.rs:79-219: integer polynomial x torus polynomial through the Fourier domain
    N = 1 << size_log
    rng = np.random.default_rng(100 + size_log)
    integer = rng.integers(-(1 << 10), 1 << 10, size=N).astype(np.int64).view(np.uint64)
    torus = rng.integers(0, 2 ** 64, size=N, dtype=np.uint64)
    fa = orc.fft_forward_as_integer(integer)
    fb = orc.fft_forward_as_torus(torus)
    out = np.zeros(N, dtype=np.uint64)
    orc.fft_add_backward_as_torus(out, fa * fb)
    ref = orc.negacyclic_mul_schoolbook(torus, integer)
    d = (out - ref).view(np.int64)
    assert np.abs(d).max() < (1 << (64 - 53 + 10 + size_log + 4))


# ----------------------------------------------------------------------------- keyswitch / PBS
def _py_keyswitch(ct, ksk, n_in, n_out, base_log, level):
    """Pure-Python restatement of lwe_keyswitch.rs:138-228 (small cases only)."""
    out = [0] * (n_out + 1)
    out[n_out] = int(ct[n_in])
    shift = 64 - base_log * level
    mask = (1 << base_log) - 1
    for i in range(n_in):
        x = int(ct[i])
        r = ((((x >> (shift - 1)) + 1) & ~1) << (shift - 1)) & M64
        state = r >> shift
        for j in range(level):  # memory index j <-> level l-j; digits come out level l first
            res = state & mask
            state >>= base_log
            carry = (((res - 1) & M64) | state) & res
            carry >>= base_log - 1
            state += carry
            digit = (res - (carry << base_log)) & M64
            row = ksk[(i * level + j) * (n_out + 1):(i * level + j + 1) * (n_out + 1)]
            for o in range(n_out + 1):
                out[o] = (out[o] - int(row[o]) * digit) & M64
    return out


def test_keyswitch_matches_pure_python_restatement():
    rng = np.random.default_rng(8)
    for (n_in, n_out, b, l) in ((9, 5, 3, 5), (16, 7, 4, 3), (4, 3, 7, 8), (5, 2, 1, 1)):
        ksk = rng.integers(0, 2 ** 64, size=n_in * l * (n_out + 1), dtype=np.uint64)
        cts = rng.integers(0, 2 ** 64, size=(6, n_in + 1), dtype=np.uint64)
        cts[0, :] = 0
        cts[1, :] = np.uint64(1) << np.uint64(63)
        cts[2, :] = np.uint64(M64)
        cts[3, :] = np.uint64(1) << np.uint64(64 - b * l - 1)
        got = orc.keyswitch(cts, ksk, n_in, n_out, b, l)
        for s in range(len(cts)):
            assert got[s].tolist() == _py_keyswitch(cts[s], ksk, n_in, n_out, b, l)


@pytest.mark.parametrize("name", ["TOY512", "TOY256K2", "TOY1024"])
def test_keyswitch_then_pbs_decrypts(name):
    ks = KeySet(PARAMS[name], seed=42)
    p = ks.p
    bits = min(p.msg_bits, 3)
    msgs = (np.arange(12) % (1 << bits)).astype(np.uint64)
    cts = ks.encrypt_big(msgs, bits)
    assert np.array_equal(ks.decrypt_big(cts, bits), msgs)
    mid = orc.keyswitch(cts, ks.ksk, p.big_n, p.n, p.b_ks, p.l_ks)
    assert np.array_equal(ks.decrypt_small(mid, bits), msgs)
    table = ((np.arange(1 << bits) * 3 + 1) % (1 << bits)).astype(np.uint64)
    lut = trivial_glwe(expand_lut(table, p.N, bits), p.k)
    out = orc.bootstrap(mid, lut, None, ks.fbsk, p.n, p.k, p.N, p.b, p.l)
    assert np.array_equal(ks.decrypt_big(out, bits), table[msgs])
    # blind rotation + separate sample extraction == bootstrap
    acc = orc.blind_rotate(mid, lut, None, ks.fbsk, p.n, p.k, p.N, p.b, p.l)
    for s in range(len(msgs)):
        assert np.array_equal(orc.sample_extract(acc[s], p.k, p.N, 0), out[s])


def test_pbs_secure_parameter_set_decrypts():
    """TEST_PARAMS_4_BITS_NATIVE_U64 (algorithms/test/mod.rs:56-78): n=742,k=1,N=2048,l=1,b=23."""
    ks = KeySet(PARAMS["P2048"], seed=42)
    p = ks.p
    bits = 4
    msgs = np.arange(16, dtype=np.uint64)
    mid = orc.keyswitch(ks.encrypt_big(msgs, bits), ks.ksk, p.big_n, p.n, p.b_ks, p.l_ks)
    table = (np.arange(16) ** 2 % 16).astype(np.uint64)
    lut = trivial_glwe(expand_lut(table, p.N, bits), p.k)
    out = orc.bootstrap(mid, lut, None, ks.fbsk, p.n, p.k, p.N, p.b, p.l)
    assert np.array_equal(ks.decrypt_big(out, bits), table[msgs])


def test_golden_fixture_is_reproducible():
    """tests/golden/ks_pbs_toy512.npz was written by tests/golden/make_golden.py; the oracle must
    reproduce its keyswitch words exactly and its PBS decryptions."""
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "ks_pbs_toy512.npz"))
    ks = KeySet(PARAMS["TOY512"], seed=42)
    p = ks.p
    assert np.array_equal(orc.keyswitch(g["cts"], ks.ksk, p.big_n, p.n, p.b_ks, p.l_ks),
                          g["ks_out"])
    out = orc.bootstrap(g["ks_out"], g["lut"], None, ks.fbsk, p.n, p.k, p.N, p.b, p.l)
    assert np.array_equal(out, g["pbs_out"])  # same binary, same rounding: deterministic
    assert np.array_equal(ks.decrypt_big(out, 3), g["expected"])


def test_encode_decode_and_lut_expansion():
    msgs = np.arange(16, dtype=np.uint64)
    assert np.array_equal(decode(encode(msgs, 4), 4), msgs)
    lut = expand_lut(np.arange(4, dtype=np.uint64), 32, 2)
    # wrappers.cpp:388-450: half a mega-case of table[0] first, negated table[0] at the end
    assert lut[0] == 0 and lut[4] == np.uint64(1) << np.uint64(61)
    assert lut[3] == 0 and lut[12] == np.uint64(2) << np.uint64(61)


@pytest.mark.parametrize("name,bits,num", [("TOY512", 3, 21), ("TOY256K2", 2, 9), ("TOY1024", 3, 17)])
def test_timed_cpu_baseline_agrees_with_the_oracle(name, bits, num):
    """oracle/tfhe_fast.cpp (the vectorised CPU implementation bench.py TIMES) against the scalar
    oracle (the CHECKER): keyswitch words equal, PBS decryptions equal, PBS phases within the
    tolerance two f64 implementations can be held to (harness.phase_tolerance).  Includes a zero
    mask element (skipped CMUX) and a ragged last SIMD batch."""
    from harness import phase_distance, phase_tolerance

    ks = KeySet(PARAMS[name], seed=42)
    p = ks.p
    msgs = (np.arange(num) % (1 << bits)).astype(np.uint64)
    cts = ks.encrypt_big(msgs, bits, seed=3)
    ref_mid = orc.keyswitch(cts, ks.ksk, p.big_n, p.n, p.b_ks, p.l_ks)
    assert np.array_equal(orc.fast_keyswitch(cts, ks.ksk, p.big_n, p.n, p.b_ks, p.l_ks), ref_mid)
    ref_mid[2, 1] = 0
    table = ((np.arange(1 << bits) * 3 + 1) % (1 << bits)).astype(np.uint64)
    lut = trivial_glwe(expand_lut(table, p.N, bits), p.k)
    ref = orc.bootstrap(ref_mid, lut, None, ks.fbsk, p.n, p.k, p.N, p.b, p.l)
    got = orc.fast_bootstrap(ref_mid, lut, None, ks.fbsk, p.n, p.k, p.N, p.b, p.l)
    assert np.array_equal(ks.decrypt_big(got, bits), ks.decrypt_big(ref, bits))
    assert phase_distance(ks, got, ref).max() <= phase_tolerance(p)


# ------------------------------------------------------------------ seeded keys
def test_aes128_fips197_vectors():
    """The reference's own known-answer test (concrete-csprng/src/generators/implem/soft/
    block_cipher.rs:90-112: FIPS-197 appendix C.1) and the appendix B example."""
    assert orc.aes128_encrypt(bytes(range(16)), bytes.fromhex("00112233445566778899aabbccddeeff")) \
        == bytes.fromhex("69c4e0d86a7b0430d8cdb78070b4c55a")
    assert orc.aes128_encrypt(bytes.fromhex("2b7e151628aed2a6abf7158809cf4f3c"),
                              bytes.fromhex("3243f6a8885a308d313198a2e0370734")) \
        == bytes.fromhex("3925841d02dc09fbdc118597196a0b32")


def test_csprng_stream_definition():
    """Byte q of a fresh generator is byte (1 + q) of the table AES_seed(counter) with key and
    counter as little-endian u128 (block_cipher.rs:16-36, generic.rs:60-72, index.rs:28-31),
    checked against an independent AES (the `cryptography` package) when it is importable; and
    forked children read consecutive ranges of the parent's table (generic.rs:124-170,
    aes_ctr_generic_test::prop_fork): any split of a range equals the range."""
    seed = 0x0123456789ABCDEF_FEDCBA9876543210
    got = orc.csprng_bytes(seed, 0, 1000).tobytes()
    try:
        from cryptography.hazmat.primitives.ciphers import Cipher, algorithms, modes
        enc = Cipher(algorithms.AES(seed.to_bytes(16, "little")), modes.ECB()).encryptor()
        table = b"".join(enc.update(i.to_bytes(16, "little")) for i in range(64))
        assert got == table[1:1001]
    except ImportError:
        pass
    block0 = orc.aes128_encrypt(seed.to_bytes(16, "little"), (0).to_bytes(16, "little"))
    assert got[:15] == block0[1:]
    for a, b in ((1, 14), (15, 1), (16, 17), (333, 500)):
        assert orc.csprng_bytes(seed, a, b).tobytes() == got[a:a + b]
    words = orc.csprng_u64(seed, 5, 40)
    assert [int(w) for w in words] == [int.from_bytes(got[8 * (5 + i): 8 * (6 + i)], "little")
                                       for i in range(40)]


@pytest.mark.parametrize("name", ["TOY512", "TOY256K2"])
def test_seeded_key_decompression_roundtrip(name):
    """generate_seeded_* / decompress_seeded_* (c_api/bootstrap.rs:120-237): a key generated with
    the compression generator's masks, stripped to its bodies and decompressed again is the same
    key word for word, masks in generator order (GGSW, level, row, polynomial); and it works: a
    keyswitch + PBS through the decompressed keys decrypts."""
    from harness import KeySet, PARAMS, expand_lut, trivial_glwe
    p = PARAMS[name]
    ks = KeySet(p, seed=42, fourier_oracle=False, compression_seeds=(0xA5A5_0000_1234, 2 ** 127 + 99))
    rows = p.n * p.l * (p.k + 1)
    full = ks.bsk.reshape(rows, p.k + 1, p.N)
    assert np.array_equal(full[:, :p.k].reshape(-1), orc.csprng_u64(ks.bsk_seed, 0, rows * p.k * p.N))
    assert np.array_equal(orc.decompress_seeded_bsk(ks.seeded_bsk, p.n, p.k, p.N, p.l, ks.bsk_seed),
                          ks.bsk)
    assert np.array_equal(orc.decompress_seeded_ksk(ks.seeded_ksk, p.big_n, p.n, p.l_ks, ks.ksk_seed),
                          ks.ksk)
    bits, num = 2, 9
    msgs = (np.arange(num) % (1 << bits)).astype(np.uint64)
    table = (np.arange(1 << bits) * 3 % (1 << bits)).astype(np.uint64)
    lut = trivial_glwe(expand_lut(table, p.N, bits), p.k)
    mid = orc.keyswitch(ks.encrypt_big(msgs, bits), ks.ksk, p.big_n, p.n, p.b_ks, p.l_ks)
    fbsk = orc.bsk_to_fourier(ks.bsk, p.n, p.k, p.l, p.N)
    out = orc.bootstrap(mid, lut, None, fbsk, p.n, p.k, p.N, p.b, p.l)
    assert np.array_equal(ks.decrypt_big(out, bits), table[msgs])
