"""GPU parity tests for seeded-key decompression on the device (csrc/seeded.cu).

Oracle: oracle/csprng_oracle.cpp — the reference's AES-128-CTR generator and its
decompress_seeded_lwe_{bootstrap,keyswitch}_key loops restated on the CPU
(backends/concrete-cpu/implementation/src/c_api/bootstrap.rs:169-237).  Bar: every word of the
decompressed key equal; the Fourier key built from the seeded form equal bit for bit to the one
built from the full key (same conversion kernel), hence identical PBS outputs."""
import numpy as np
import pytest

from harness import KeySet, PARAMS, expand_lut, orc, trivial_glwe

pytestmark = pytest.mark.gpu

SEEDS = (0x0123456789ABCDEF_FEDCBA9876543210, 2 ** 127 + 0xDEADBEEF)


@pytest.mark.parametrize("name,n_small", [("TOY512", None), ("TOY256K2", None), ("P2048", 64),
                                          ("P4", 40), ("P8", 6)])
def test_seeded_bsk_decompression_bit_exact(engine, name, n_small):
    """k = 1, 2; l = 1, 2, 3; N = 256 ... 16384 — masks in generator order, bodies in place."""
    from concrete_b200 import as_u64

    ks = KeySet(PARAMS[name], seed=42, fourier_oracle=False, n_override=n_small,
                compression_seeds=SEEDS)
    p = ks.p
    ref = orc.decompress_seeded_bsk(ks.seeded_bsk, p.n, p.k, p.N, p.l, ks.bsk_seed)
    assert np.array_equal(ref, ks.bsk)
    got = as_u64(engine.decompress_seeded_bsk(ks.seeded_bsk, p.n, p.k, p.l, p.N, ks.bsk_seed))
    assert np.array_equal(got, ref)


@pytest.mark.parametrize("n_in,n_out,level", [(512, 32, 4), (1024, 743, 5), (2048, 742, 5),
                                              (37, 5, 1), (3, 1000, 2)])
def test_seeded_ksk_decompression_bit_exact(engine, n_in, n_out, level):
    """Mask rows of n_out words (not a multiple of the 8 words a thread produces), body last."""
    from concrete_b200 import as_u64

    rng = np.random.default_rng(n_in + n_out)
    seeded = rng.integers(0, 2 ** 64, size=n_in * level, dtype=np.uint64)
    seed = int(rng.integers(0, 2 ** 63)) << 64 | int(rng.integers(0, 2 ** 63))
    ref = orc.decompress_seeded_ksk(seeded, n_in, n_out, level, seed)
    got = as_u64(engine.decompress_seeded_ksk(seeded, n_in, n_out, level, seed))
    assert np.array_equal(got, ref)
    # bodies from a DEVICE pointer as well
    import torch
    dest = torch.empty(ref.size, dtype=torch.int64, device=engine.device)
    d_seeded = engine.to_device(seeded)
    engine.lib.cuda_decompress_seeded_lwe_keyswitch_key_64(
        engine.stream, engine.gpu, dest.data_ptr(), d_seeded.data_ptr(), n_in, n_out, level,
        seed & (2 ** 64 - 1), seed >> 64)
    engine.synchronize()
    assert np.array_equal(as_u64(dest), ref)


def test_seeded_empty_key(engine):
    import torch
    dest = torch.zeros(8, dtype=torch.int64, device=engine.device)
    engine.lib.cuda_decompress_seeded_lwe_keyswitch_key_64(
        engine.stream, engine.gpu, dest.data_ptr(), dest.data_ptr(), 0, 5, 3, 1, 2)
    engine.synchronize()
    assert int(dest.abs().sum()) == 0


@pytest.mark.parametrize("name,n_small,bits", [("TOY1024", None, 3), ("P2048", 24, 4), ("P7", 10, 4),
                                               ("P8", 6, 4)])
def test_seeded_keys_drive_the_hot_path(engine, name, n_small, bits):
    """Keyswitch + PBS with both keys ingested in seeded form (masks regenerated on the device,
    BSK converted to the Fourier domain in place): keyswitch words equal the oracle's, the Fourier
    key equals the one converted from the full key bit for bit, so the PBS outputs are identical,
    and every decryption is LUT[m]."""
    from concrete_b200 import as_u64

    ks = KeySet(PARAMS[name], seed=42, n_override=n_small, compression_seeds=SEEDS)
    p = ks.p
    num = 21
    rng = np.random.default_rng(5)
    msgs = rng.integers(0, 1 << bits, num).astype(np.uint64)
    table = ((np.arange(1 << bits) * 5 + 3) % (1 << bits)).astype(np.uint64)
    lut = trivial_glwe(expand_lut(table, p.N, bits), p.k)
    cts = ks.encrypt_big(msgs, bits, seed=17)
    d_ksk = engine.decompress_seeded_ksk(ks.seeded_ksk, p.big_n, p.n, p.l_ks, ks.ksk_seed)
    d_fbsk = engine.decompress_seeded_bsk(ks.seeded_bsk, p.n, p.k, p.l, p.N, ks.bsk_seed,
                                          convert=True)
    d_fbsk_full = engine.convert_bsk(ks.bsk, p.n, p.k, p.l, p.N)
    assert np.array_equal(as_u64(d_fbsk), as_u64(d_fbsk_full))
    d_mid = engine.keyswitch(engine.to_device(cts), d_ksk, p.big_n, p.n, p.b_ks, p.l_ks)
    assert np.array_equal(as_u64(d_mid), orc.keyswitch(cts, ks.ksk, p.big_n, p.n, p.b_ks, p.l_ks))
    d_lut = engine.to_device(lut)
    out = as_u64(engine.bootstrap(d_mid, d_lut, None, d_fbsk, p.n, p.k, p.N, p.b, p.l))
    out_full = as_u64(engine.bootstrap(d_mid, d_lut, None, d_fbsk_full, p.n, p.k, p.N, p.b, p.l))
    assert np.array_equal(out, out_full)
    assert np.array_equal(ks.decrypt_big(out, bits), table[msgs])


def test_seeded_keys_through_the_runtime_wrappers(engine):
    """memref_batched_* with the context holding SEEDED keys (what the keyset ships): the
    per-device key copies are built by the device-side decompression on first use; outputs equal
    those of a context holding the full keys, word for word."""
    from concrete_b200.runtime import Runtime

    ks = KeySet(PARAMS["TOY1024"], seed=42, compression_seeds=SEEDS)
    p = ks.p
    bits, num = 3, 40
    rng = np.random.default_rng(9)
    msgs = rng.integers(0, 1 << bits, num).astype(np.uint64)
    table = (np.arange(1 << bits) ** 2 % (1 << bits)).astype(np.uint64)
    lut_poly = expand_lut(table, p.N, bits)
    cts = ks.encrypt_big(msgs, bits, seed=3)
    outs = []
    for seeded in (True, False):
        rt = Runtime(num_gpus=1)
        if seeded:
            ki = rt.add_seeded_ksk(ks.seeded_ksk, ks.ksk.size, ks.ksk_seed)
            bi = rt.add_seeded_bsk(ks.seeded_bsk, ks.bsk.size, ks.bsk_seed)
        else:
            ki, bi = rt.add_ksk(ks.ksk), rt.add_bsk(ks.bsk)
        mid = np.zeros((num, p.n + 1), dtype=np.uint64)
        out = np.zeros((num, p.big_n + 1), dtype=np.uint64)
        rt.batched_keyswitch(mid, cts, p.l_ks, p.b_ks, p.big_n, p.n, ki)
        rt.batched_bootstrap(out, mid, lut_poly, p.n, p.N, p.l, p.b, p.k, bi)
        rt.close()
        outs.append((mid, out))
    assert np.array_equal(outs[0][0], orc.keyswitch(cts, ks.ksk, p.big_n, p.n, p.b_ks, p.l_ks))
    assert np.array_equal(outs[0][0], outs[1][0]) and np.array_equal(outs[0][1], outs[1][1])
    assert np.array_equal(ks.decrypt_big(outs[0][1], bits), table[msgs])
