"""GPU parity tests for WoP-PBS on the device (csrc/wop_pbs.cu).

Oracle: oracle/wop_oracle.inl — extract_bits, circuit_bootstrap_boolean, the CMUX tree, the GGSW
blind rotation and memref_wop_pbs_crt_buffer restated on the CPU
(tfhe-rs core_crypto/fft_impl/fft64/crypto/wop_pbs/mod.rs, compilers/concrete-compiler/compiler/
lib/Runtime/wrappers.cpp:855-997).  Bars: integer-only results bit-exact (the first extracted bit
is a pure keyswitch); everything behind an f64 PBS by decryption — every extracted bit, every CRT
residue of LUT[m] — with the output noise held to a fraction of the encoding step, for the GPU
and for the oracle on the same inputs (the reference's own tests pin this path the same way:
wop_pbs/tests.rs:94-250,580-760)."""
import numpy as np
import pytest

from harness import orc
from wop_harness import (WOP_PARAMS, WopKeySet, crt_bits, crt_decode, crt_encode,
                         encode_lut_for_crt_woppbs)

pytestmark = pytest.mark.gpu

_KS = {}


def wop_keys(name):
    if name not in _KS:
        _KS[name] = WopKeySet(WOP_PARAMS[name])
    return _KS[name]


def device_keys(engine, ks):
    p = ks.p
    return (engine.convert_bsk(ks.bsk, p.n, p.k, p.bsk_l, p.N), engine.upload_ksk(ks.ksk),
            engine.to_device(ks.pfpksk))


@pytest.mark.parametrize("name,nbits", [("WOP256", [5]), ("WOP256", [3, 1, 4]), ("WOP256K2", [4, 2]),
                                        ("WOP1024", [6, 3])])
def test_extract_bits(engine, name, nbits):
    """extract_bits (wop_pbs/mod.rs:63-222) on several blocks at once: bit lists laid out last
    block first, MSB first; the least significant bit of every block is a pure keyswitch of the
    shifted input (bit-exact); every bit decrypts to the message bit on the padding position."""
    from concrete_b200 import as_u64

    ks = wop_keys(name)
    p = ks.p
    rng = np.random.default_rng(7)
    msgs = [int(rng.integers(0, 1 << nb)) for nb in nbits]
    # message in the top nb bits (delta_log = 64 - nb), plus half a step below for the rounding
    pts = [np.uint64((m << (64 - nb)) & (2 ** 64 - 1)) for m, nb in zip(msgs, nbits)]
    cts = ks.encrypt_big_raw(pts, seed=11)
    d_fbsk, d_ksk, _ = device_keys(engine, ks)
    got = as_u64(engine.extract_bits(engine.to_device(cts), nbits, d_fbsk, d_ksk, p))
    off = 0
    for i in range(len(nbits) - 1, -1, -1):
        nb = nbits[i]
        ref = orc.extract_bits(cts[i], ks.fbsk, ks.ksk, p.n, p.k, p.N, p.bsk_b, p.bsk_l, p.ks_b, p.ks_l,
                               64 - nb, nb)
        mine = got[off:off + nb]
        assert np.array_equal(mine[nb - 1], ref[nb - 1])          # LSB: integer arithmetic only
        for rows in (mine, ref):
            dec = ks.decrypt_small_raw(rows)
            bits = ((dec.astype(np.uint64) + np.uint64(1 << 62)) >> np.uint64(63)).astype(int)
            assert list(bits) == [(msgs[i] >> (nb - 1 - j)) & 1 for j in range(nb)]
            err = (dec - (bits.astype(np.uint64) << np.uint64(63))).view(np.int64)
            assert np.abs(err).max() < 2 ** 59                     # far from the decision threshold
        off += nb


@pytest.mark.parametrize("name,nbits,lut_count", [("WOP256", 5, 2), ("WOP256", 9, 1), ("WOP256", 10, 3),
                                                  ("WOP256K2", 9, 2), ("WOP1024", 6, 2),
                                                  ("WOP1024", 11, 1)])
def test_circuit_bootstrap_vertical_packing(engine, name, nbits, lut_count):
    """Circuit bootstrap of `nbits` boolean ciphertexts + vertical packing of `lut_count` tables of
    2^nbits entries: tables shorter than a polynomial (blind rotation only), exactly one polynomial,
    and longer (CMUX tree of 1-2 layers + blind rotation); k = 1 and 2."""
    from concrete_b200 import as_u64

    ks = wop_keys(name)
    p = ks.p
    rng = np.random.default_rng(nbits)
    out_bits = 4
    luts = (rng.integers(0, 1 << out_bits, size=(lut_count, 1 << nbits)).astype(np.uint64)
            << np.uint64(64 - out_bits - 1))
    d_fbsk, _, d_pfpksk = device_keys(engine, ks)
    for trial in range(3):
        value = int(rng.integers(0, 1 << nbits))
        bits = np.array([(value >> (nbits - 1 - j)) & 1 for j in range(nbits)], dtype=np.uint64)
        cts = orc.lwe_encrypt(ks.sk_small, bits << np.uint64(63), p.sigma_lwe, 50 + trial)
        ref = orc.cbs_vertical_packing(cts, luts, ks.fbsk, ks.pfpksk, p.n, p.k, p.N, p.bsk_b, p.bsk_l,
                                       p.pks_b, p.pks_l, p.cbs_b, p.cbs_l)
        got = as_u64(engine.circuit_bootstrap_vertical_packing(
            engine.to_device(cts), engine.to_device(luts), d_fbsk, d_pfpksk, p))
        expect = luts[:, value]
        for out in (got, ref):
            err = (ks.decrypt_big_raw(out) - expect).view(np.int64)
            assert np.abs(err).max() < 2 ** (64 - out_bits - 3), (name, nbits, np.log2(np.abs(err).max() + 1))
        diff = (ks.decrypt_big_raw(got) - ks.decrypt_big_raw(ref)).view(np.int64)
        assert np.abs(diff).max() < 2 ** (64 - out_bits - 3)


@pytest.mark.parametrize("name,moduli", [("WOP256", [3, 4]), ("WOP256", [7, 8, 5]), ("WOP256K2", [5, 4]),
                                         ("WOP1024", [7, 8]), ("WOP1024", [16, 128])])
def test_wop_pbs_crt(engine, name, moduli):
    """memref_wop_pbs_crt_buffer end to end (wrappers.cpp:855-997): a value held as CRT residues,
    one ciphertext per modulus, goes through bit extraction, circuit bootstrap and vertical packing
    with the table the runtime builds (memref_encode_lut_for_crt_woppbs); every residue of
    LUT[m] comes out, on the GPU and in the oracle."""
    from concrete_b200 import as_u64

    ks = wop_keys(name)
    p = ks.p
    product = int(np.prod(moduli))
    table = [(3 * v * v + 1) % product for v in range(product)]
    lut = encode_lut_for_crt_woppbs(table, moduli)
    d_fbsk, d_ksk, d_pfpksk = device_keys(engine, ks)
    d_lut = engine.to_device(lut)
    rng = np.random.default_rng(product)
    for m in [0, product - 1] + [int(v) for v in rng.integers(0, product, 3)]:
        pts = [crt_encode(m, b, product) for b in moduli]
        cts = ks.encrypt_big_raw(pts, seed=100 + m)
        got = as_u64(engine.wop_pbs_crt(engine.to_device(cts), d_lut, moduli, d_fbsk, d_ksk, d_pfpksk, p))
        ref = orc.wop_pbs_crt(cts, lut, moduli, ks.fbsk, ks.ksk, ks.pfpksk, p.n, p.k, p.N, p.bsk_b,
                              p.bsk_l, p.ks_b, p.ks_l, p.pks_b, p.pks_l, p.cbs_b, p.cbs_l)
        want = [table[m] % b for b in moduli]
        for out in (got, ref):
            dec = ks.decrypt_big_raw(out)
            assert [crt_decode(int(d), b) for d, b in zip(dec, moduli)] == want, (name, moduli, m)


def test_wop_pbs_crt_16_bit_parameters(engine):
    """The optimizer's 128-bit parameter set for 16 bits of precision through WoP-PBS
    (wop_pbs_last_128: k=2, N=1024, n=707, br 4x9, ks 6x2, cb 2x9, pp 2x16) with real noise: a
    16-bit value as residues mod 7, 8, 9, 11, 13 (18 bits in all: CMUX tree of 8 layers over five
    tables of 2^18 entries + blind rotation).  GPU and oracle both return every residue of LUT[v]."""
    from concrete_b200 import as_u64

    ks = wop_keys("WOP16")
    p = ks.p
    moduli = [7, 8, 9, 11, 13]
    product = int(np.prod(moduli))
    table = ((np.arange(1 << 16, dtype=np.int64) * 40503 + 7) & 0xFFFF)
    lut = encode_lut_for_crt_woppbs(table, moduli)
    d_fbsk, d_ksk, d_pfpksk = device_keys(engine, ks)
    d_lut = engine.to_device(lut)
    for trial, v in enumerate((0xBEEF, 65535, 1)):
        cts = ks.encrypt_big_raw([crt_encode(v, b, product) for b in moduli], seed=900 + trial)
        got = as_u64(engine.wop_pbs_crt(engine.to_device(cts), d_lut, moduli, d_fbsk, d_ksk, d_pfpksk, p))
        want = [int(table[v]) % b for b in moduli]
        dec = ks.decrypt_big_raw(got)
        assert [crt_decode(int(d), b) for d, b in zip(dec, moduli)] == want
        if trial == 0:  # the oracle at full size takes a few seconds: once
            ref = orc.wop_pbs_crt(cts, lut, moduli, ks.fbsk, ks.ksk, ks.pfpksk, p.n, p.k, p.N, p.bsk_b,
                                  p.bsk_l, p.ks_b, p.ks_l, p.pks_b, p.pks_l, p.cbs_b, p.cbs_l)
            dref = ks.decrypt_big_raw(ref)
            assert [crt_decode(int(d), b) for d, b in zip(dref, moduli)] == want


def test_wop_pbs_crt_through_the_runtime_wrapper(engine):
    """memref_wop_pbs_crt_buffer with HOST memrefs and keys held by the RuntimeContext — the
    reference's prototype (wrappers.cpp:855-884), executed on the device."""
    from concrete_b200.runtime import Runtime

    ks = wop_keys("WOP256")
    p = ks.p
    moduli = [7, 8, 5]
    product = int(np.prod(moduli))
    table = [(v * 5 + 2) % product for v in range(product)]
    lut = encode_lut_for_crt_woppbs(table, moduli)
    rt = Runtime(num_gpus=1)
    ki, bi, pi = rt.add_ksk(ks.ksk), rt.add_bsk(ks.bsk), rt.add_pksk(ks.pfpksk)
    for m in (1, 123, product - 1):
        cts = ks.encrypt_big_raw([crt_encode(m, b, product) for b in moduli], seed=m)
        out = np.zeros_like(cts)
        rt.wop_pbs_crt(out, cts, lut, moduli, p, ki, bi, pi)
        dec = ks.decrypt_big_raw(out)
        assert [crt_decode(int(d), b) for d, b in zip(dec, moduli)] == [table[m] % b for b in moduli]
    rt.close()
