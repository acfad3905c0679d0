"""GPU parity tests for the composite workloads of BASELINE.json `configs`:

  cfg 1  4-bit table lookup x**2 on an encrypted uint4 tensor of 64 elements (P4 parameters),
         through the runtime wrappers with host buffers, as the compiled program calls them;
  cfg 4  Concrete-ML-style dense layer (clear weights x encrypted activations, leveled ops on the
         device) followed by a batch of ReLU table lookups (P8's polynomial size, N = 16384);
  cfg 5  16-bit addition on values stored as 8 chunks of 2 bits inside 4-bit ciphertexts
         (chunkSize 4, chunkWidth 2: compilers/concrete-compiler/compiler/lib/Dialect/FHE/
         Transforms/BigInt.cpp:55-60): ripple carry, one carry-extract table lookup per chunk =
         8 strictly dependent keyswitch+PBS rounds (BigInt.cpp:109-154), then one per-chunk TLU.

Every device step goes through the C ABI.  Checker: the oracle runs the same op sequence on the
same ciphertext bytes — leveled ops and keyswitch outputs are compared word for word, PBS
outputs by decryption (and the final decrypted values against the plaintext computation)."""
import numpy as np
import pytest

from harness import KeySet, PARAMS, expand_lut, orc, trivial_glwe

pytestmark = pytest.mark.gpu


def _dev_add(engine, a, b, dim):
    import torch
    out = torch.empty_like(a)
    engine.lib.cuda_add_lwe_ciphertext_vector_64(engine.stream, engine.gpu, out.data_ptr(),
                                                 a.data_ptr(), b.data_ptr(), dim, a.shape[0])
    return out


def _dev_mul_clear(engine, a, clear_u64, dim):
    import torch
    out = torch.empty_like(a)
    d_c = engine.to_device(np.ascontiguousarray(clear_u64, dtype=np.uint64))
    engine.lib.cuda_mult_lwe_ciphertext_vector_cleartext_vector_64(
        engine.stream, engine.gpu, out.data_ptr(), a.data_ptr(), d_c.data_ptr(), dim, a.shape[0])
    return out


def _dev_add_plain(engine, a, plain_u64, dim):
    import torch
    out = torch.empty_like(a)
    d_p = engine.to_device(np.ascontiguousarray(plain_u64, dtype=np.uint64))
    engine.lib.cuda_add_lwe_ciphertext_vector_plaintext_vector_64(
        engine.stream, engine.gpu, out.data_ptr(), a.data_ptr(), d_p.data_ptr(), dim, a.shape[0])
    return out


def test_cfg1_uint4_square_tensor64(engine):
    """configs[0]: x**2 mod 16 on 64 encrypted uint4, optimizer parameters for 4 bits (P4)."""
    from concrete_b200.runtime import Runtime
    from harness import phase_distance, phase_tolerance

    p = PARAMS["P4"]
    ks = KeySet(p, seed=42)
    bits, num = 4, 64
    table = (np.arange(16) ** 2 % 16).astype(np.uint64)
    lut_poly = expand_lut(table, p.N, bits)
    msgs = (np.arange(num) % 16).astype(np.uint64)
    cts = ks.encrypt_big(msgs, bits)
    rt = Runtime(num_gpus=1)
    ki, bi = rt.add_ksk(ks.ksk), rt.add_bsk(ks.bsk)
    mid = np.zeros((num, p.n + 1), dtype=np.uint64)
    out = np.zeros((num, p.big_n + 1), dtype=np.uint64)
    rt.batched_keyswitch(mid, cts, p.l_ks, p.b_ks, p.big_n, p.n, ki)
    rt.batched_bootstrap(out, mid, lut_poly, p.n, p.N, p.l, p.b, p.k, bi)
    rt.close()
    ref_mid = orc.keyswitch(cts, ks.ksk, p.big_n, p.n, p.b_ks, p.l_ks)
    ref_out = orc.bootstrap(ref_mid, trivial_glwe(lut_poly, p.k), None, ks.fbsk, p.n, p.k, p.N,
                            p.b, p.l)
    assert np.array_equal(mid, ref_mid)
    assert np.array_equal(ks.decrypt_big(out, bits), table[msgs])
    assert np.array_equal(ks.decrypt_big(ref_out, bits), table[msgs])
    assert phase_distance(ks, out, ref_out).max() <= phase_tolerance(p)


def test_cfg4_dense_layer_relu(engine):
    """configs[3]: z = W x + b on ciphertexts (leveled, device), then ReLU by table lookup on the
    whole layer at once.  N = 16384 / l = 2 as in P8; the small dimension is cut to 40 so that the
    oracle finishes in seconds (key layouts and kernels are those of the full set)."""
    import torch
    from concrete_b200 import as_u64

    ks = KeySet(PARAMS["P8"], seed=5, n_override=40)
    p = ks.p
    bits, n_in, n_out, batch = 6, 16, 24, 5
    rng = np.random.default_rng(8)
    x = rng.integers(0, 4, size=(batch, n_in)).astype(np.int64)
    W = np.zeros((n_out, n_in), dtype=np.int64)           # 8 weights of +-1 per neuron: |W x| <= 24
    for j in range(n_out):
        W[j, rng.choice(n_in, 8, replace=False)] = rng.choice([-1, 1], 8)
    bias = rng.integers(-6, 7, size=n_out).astype(np.int64)
    z = x @ W.T + bias
    assert np.abs(z).max() < 32                           # 6-bit signed
    relu = np.maximum(z, 0)
    # signed values travel with an offset of 2^(bits-1) (the padding bit must stay clear)
    table = np.array([max(v - 32, 0) for v in range(64)], dtype=np.uint64)
    lut = trivial_glwe(expand_lut(table, p.N, bits), p.k)

    cts = ks.encrypt_big(x.reshape(-1).astype(np.uint64), bits).reshape(batch, n_in, p.big_n + 1)
    d_cts = engine.to_device(cts.reshape(batch * n_in, -1))
    # dense layer on the device: for every output neuron, ct_i * w_ji summed over i, + (b_j + 32)
    shift = np.uint64(63 - bits)
    pre_rows, ref_rows = [], []
    for j in range(n_out):
        w = np.tile(W[j].astype(np.uint64), batch)        # wrapping u64 clear multipliers
        prod = _dev_mul_clear(engine, d_cts, w, p.big_n).reshape(batch, n_in, -1)
        acc = prod[:, 0].contiguous()
        for i in range(1, n_in):
            acc = _dev_add(engine, acc, prod[:, i].contiguous(), p.big_n)
        off = np.full(batch, np.uint64((int(bias[j]) + 32) % (1 << 64)) << shift, dtype=np.uint64)
        pre_rows.append(_dev_add_plain(engine, acc, off, p.big_n))
        # oracle side: the same wrapping arithmetic in numpy
        r = (cts * W[j].astype(np.uint64)[None, :, None]).sum(axis=1, dtype=np.uint64)
        r[:, p.big_n] += off
        ref_rows.append(r)
    d_pre = torch.stack(pre_rows, dim=1).reshape(batch * n_out, -1).contiguous()
    ref_pre = np.stack(ref_rows, axis=1).reshape(batch * n_out, -1)
    assert np.array_equal(as_u64(d_pre), ref_pre)          # leveled ops: bit-exact
    # ReLU TLUs: keyswitch + PBS over the whole layer
    d_mid = engine.keyswitch(d_pre, engine.upload_ksk(ks.ksk), p.big_n, p.n, p.b_ks, p.l_ks)
    ref_mid = orc.keyswitch(ref_pre, ks.ksk, p.big_n, p.n, p.b_ks, p.l_ks)
    assert np.array_equal(as_u64(d_mid), ref_mid)
    d_fbsk = engine.convert_bsk(ks.bsk, p.n, p.k, p.l, p.N)
    d_out = engine.bootstrap(d_mid, engine.to_device(lut), None, d_fbsk, p.n, p.k, p.N, p.b, p.l)
    ref_out = orc.bootstrap(ref_mid, lut, None, ks.fbsk, p.n, p.k, p.N, p.b, p.l)
    got = ks.decrypt_big(as_u64(d_out), bits).reshape(batch, n_out)
    assert np.array_equal(got, relu.astype(np.uint64))
    assert np.array_equal(ks.decrypt_big(ref_out, bits).reshape(batch, n_out), got)


def test_cfg5_chunked_16bit_add(engine):
    """configs[4]: 16-bit a + b as 8 chunks of 2 bits in 4-bit ciphertexts (P4 parameters, full
    size).  Round c: s_c = a_c + b_c + carry_(c-1) (leveled), carry_c = TLU[s >> 2](s_c) — the
    dependent keyswitch+PBS — and, after the loop, one batched TLU[s & 3] over all chunks."""
    import torch
    from concrete_b200 import as_u64

    p = PARAMS["P4"]
    ks = KeySet(p, seed=42)
    bits, ops, chunks = 4, 48, 8
    rng = np.random.default_rng(16)
    a = rng.integers(0, 1 << 16, ops)
    b = rng.integers(0, 1 << 16, ops)
    a[0], b[0] = 0xFFFF, 0x0001                           # carry through every chunk
    a[1], b[1] = 0xFFFF, 0xFFFF
    split = lambda v: np.stack([(v >> (2 * c)) & 3 for c in range(chunks)], axis=1).astype(np.uint64)
    ca = ks.encrypt_big(split(a).reshape(-1), bits, seed=21).reshape(ops, chunks, -1)
    cb = ks.encrypt_big(split(b).reshape(-1), bits, seed=22).reshape(ops, chunks, -1)
    carry_tab = (np.arange(16) >> 2).astype(np.uint64)
    low_tab = (np.arange(16) & 3).astype(np.uint64)
    lut_carry = trivial_glwe(expand_lut(carry_tab, p.N, bits), p.k)
    lut_low = trivial_glwe(expand_lut(low_tab, p.N, bits), p.k)
    d_ksk = engine.upload_ksk(ks.ksk)
    d_fbsk = engine.convert_bsk(ks.bsk, p.n, p.k, p.l, p.N)
    d_lut_carry, d_lut_low = engine.to_device(lut_carry), engine.to_device(lut_low)
    d_a, d_b = engine.to_device(ca.reshape(ops * chunks, -1)), engine.to_device(cb.reshape(ops * chunks, -1))
    d_sum = _dev_add(engine, d_a, d_b, p.big_n).reshape(ops, chunks, -1)
    ref_sum = (ca + cb)
    d_carry, ref_carry = None, None
    sums, ref_sums = [], []
    for c in range(chunks):                               # 8 dependent KS+PBS rounds
        s_c, r_c = d_sum[:, c].contiguous(), np.ascontiguousarray(ref_sum[:, c])
        if d_carry is not None:
            s_c = _dev_add(engine, s_c, d_carry, p.big_n)
            r_c = r_c + ref_carry
        sums.append(s_c)
        ref_sums.append(r_c)
        d_mid = engine.keyswitch(s_c, d_ksk, p.big_n, p.n, p.b_ks, p.l_ks)
        d_carry = engine.bootstrap(d_mid, d_lut_carry, None, d_fbsk, p.n, p.k, p.N, p.b, p.l)
        # the oracle follows ITS OWN chain (PBS words differ between f64 implementations, the
        # decrypted carries must not); the keyswitch is compared on identical inputs below
        ref_mid = orc.keyswitch(r_c, ks.ksk, p.big_n, p.n, p.b_ks, p.l_ks)
        ref_carry = orc.bootstrap(ref_mid, lut_carry, None, ks.fbsk, p.n, p.k, p.N, p.b, p.l)
        assert np.array_equal(ks.decrypt_big(as_u64(d_carry), bits), ks.decrypt_big(ref_carry, bits))
        same_in = as_u64(s_c)
        assert np.array_equal(as_u64(d_mid), orc.keyswitch(same_in, ks.ksk, p.big_n, p.n, p.b_ks, p.l_ks))
    d_all = torch.stack(sums, dim=1).reshape(ops * chunks, -1).contiguous()
    d_mid = engine.keyswitch(d_all, d_ksk, p.big_n, p.n, p.b_ks, p.l_ks)
    d_low = engine.bootstrap(d_mid, d_lut_low, None, d_fbsk, p.n, p.k, p.N, p.b, p.l)
    low = ks.decrypt_big(as_u64(d_low), bits).reshape(ops, chunks).astype(np.int64)
    total = sum(low[:, c] << (2 * c) for c in range(chunks))
    assert np.array_equal(total, (a + b) & 0xFFFF)
    assert np.array_equal(ks.decrypt_big(as_u64(d_carry), bits).astype(np.int64), (a + b) >> 16)


def test_wrappers_sharded_over_two_gpus(engine):
    """the batched wrappers shard a batch over SDFG_NUM_GPUS devices (contiguous
    split, one host thread + stream per device, keys uploaded/converted per device, no collective
    in the data path).  Needs a box with >= 2 GPUs; the split rule itself is covered on CPU by
    tests/test_host_logic.py."""
    import torch
    from concrete_b200.runtime import Runtime

    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    ks = KeySet(PARAMS["TOY1024"], seed=42)
    p = ks.p
    bits, num = 3, 2 * 888 * 2 + 101                      # odd split, pipelined chunks on each GPU
    table = (np.arange(8) ** 2 % 8).astype(np.uint64)
    lut_poly = expand_lut(table, p.N, bits)
    rng = np.random.default_rng(23)
    msgs = rng.integers(0, 8, num).astype(np.uint64)
    cts = ks.encrypt_big(msgs, bits)
    ref_mid = orc.keyswitch(cts, ks.ksk, p.big_n, p.n, p.b_ks, p.l_ks)
    rt = Runtime(num_gpus=2)
    ki, bi = rt.add_ksk(ks.ksk), rt.add_bsk(ks.bsk)
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


@pytest.mark.parametrize("name,num,bits", [("P7", 1024, 6), ("P8", 512, 7), ("P6", 2048, 5)])
def test_cfg3_large_polynomial_full_size(engine, name, num, bits):
    """configs[2]: full parameter sets with N = 4096 / 8192 / 16384 (all n CMUX iterations, real
    noise, one bit of margin below the precision the set was optimised for).  Every decryption
    equals LUT[m] after keyswitch + PBS; a spot sample of keyswitch outputs matches the oracle
    word for word; and 12 samples go through the oracle's PBS at FULL n as well: same
    decryptions, every phase within the PBS tolerance (harness.phase_tolerance)."""
    from concrete_b200 import as_u64
    from harness import phase_distance, phase_tolerance

    p = PARAMS[name]
    ks = KeySet(p, seed=42, fourier_oracle=True)
    rng = np.random.default_rng(31)
    msgs = rng.integers(0, 1 << bits, num).astype(np.uint64)
    table = ((np.arange(1 << bits) * 3 + 1) % (1 << bits)).astype(np.uint64)
    lut = trivial_glwe(expand_lut(table, p.N, bits), p.k)
    cts = ks.encrypt_big(msgs, bits, seed=5)
    d_mid = engine.keyswitch(engine.to_device(cts), engine.upload_ksk(ks.ksk), p.big_n, p.n,
                             p.b_ks, p.l_ks)
    d_fbsk = engine.convert_bsk(ks.bsk, p.n, p.k, p.l, p.N)
    d_out = engine.bootstrap(d_mid, engine.to_device(lut), None, d_fbsk, p.n, p.k, p.N, p.b, p.l)
    engine.synchronize()
    got = as_u64(d_out)
    assert np.array_equal(ks.decrypt_big(got, bits), table[msgs])
    pick = np.sort(rng.choice(num, 12, replace=False))
    ref_mid = orc.keyswitch(np.ascontiguousarray(cts[pick]), ks.ksk, p.big_n, p.n, p.b_ks, p.l_ks)
    assert np.array_equal(as_u64(d_mid)[pick], ref_mid)
    ref_out = orc.bootstrap(ref_mid, lut, None, ks.fbsk, p.n, p.k, p.N, p.b, p.l)
    assert np.array_equal(ks.decrypt_big(ref_out, bits), table[msgs[pick]])
    dist = phase_distance(ks, np.ascontiguousarray(got[pick]), ref_out)
    assert dist.max() <= phase_tolerance(p), (np.log2(dist.max() + 1), np.log2(phase_tolerance(p)))


def pbs_variance_model(p):
    """Output-noise variance of one PBS in units of (2^-64)^2: n * variance_cmux with the
    reference's noise model (backends/concrete-cpu/noise-model/src/gaussian_noise/noise/
    external_product_glwe.rs:31-89 summed by blind_rotate.rs:8-28; binary keys, fft_precision 53)."""
    k, N, l = float(p.k), float(p.N), float(p.l)
    b, q2 = 2.0 ** p.b, 2.0 ** 128
    b2l = 2.0 ** (2 * p.b * p.l)
    var_ggsw = p.sigma_glwe ** 2                          # as a fraction of the torus, squared
    res1 = l * (k + 1) * N * (b * b + 2) / 12.0 * var_ggsw
    res2 = ((q2 - b2l) / (24.0 * b2l) * (1.0 + k * N * (0.25 + 0.25)) + k * N / 8.0 * 0.25
            + (1.0 - k * N) ** 2 / 16.0 * 0.25) / q2
    fft = 2.0 ** -2.57722494 * 2.0 ** (2 * (64 - 53)) * l * b * b * N * N * (k + 1) / q2
    return p.n * (res1 + res2 + fft) * q2


def test_pbs_output_noise_variance(engine):
    """The reference pins PBS quality statistically (tfhe-rs core_crypto/algorithms/test/
    noise_distribution/lwe_programmable_boostrapping_noise.rs:35-190): the measured variance of
    the output noise must not exceed the noise-model formula (plus its 10 % tolerance) and the
    noise must be centred.  1024 P2048 bootstraps on the GPU; the oracle's output on a subset of
    the same inputs is held to the same bound, so a biased FFT or rounding on either side shows."""
    from concrete_b200 import as_u64

    p = PARAMS["P2048"]
    ks = KeySet(p, seed=42)
    bits, num, num_ref = 4, 1024, 96
    rng = np.random.default_rng(77)
    msgs = rng.integers(0, 1 << bits, num).astype(np.uint64)
    table = np.arange(1 << bits, dtype=np.uint64)           # identity, as the reference test
    lut = trivial_glwe(expand_lut(table, p.N, bits), p.k)
    mid = ks.encrypt_small(msgs, bits, seed=123)
    d_fbsk = engine.convert_bsk(ks.bsk, p.n, p.k, p.l, p.N)
    got = as_u64(engine.bootstrap(engine.to_device(mid), engine.to_device(lut), None, d_fbsk, p.n,
                                  p.k, p.N, p.b, p.l))
    ref = orc.bootstrap(np.ascontiguousarray(mid[:num_ref]), lut, None, ks.fbsk, p.n, p.k, p.N,
                        p.b, p.l)
    expect = (table[msgs] << np.uint64(63 - bits)).astype(np.uint64)
    model = pbs_variance_model(p)
    for out, cnt in ((got, num), (ref, num_ref)):
        assert np.array_equal(ks.decrypt_big(out, bits), table[msgs[:cnt]])
        noise = (orc.lwe_decrypt(ks.sk_glwe, out) - expect[:cnt]).view(np.int64).astype(np.float64)
        var, mean = noise.var(), noise.mean()
        # chi-square spread of a variance estimate from cnt samples: std = var * sqrt(2/cnt)
        assert var <= model * 1.10 * (1.0 + 4.0 * np.sqrt(2.0 / cnt)), (np.log2(var), np.log2(model))
        assert var >= model / 64.0, (np.log2(var), np.log2(model))     # not trivially noiseless
        assert abs(mean) <= 4.0 * np.sqrt(var / cnt), (mean, np.sqrt(var / cnt))
