"""GPU parity tests of the SDFG ("stream emulator") entry points — the path the compiled program
takes under fhe.Compiler(..., use_gpu=True) (GPUDFG.cpp:1467-1799; API
stream_emulator_api.h:29-104).  Host memrefs in, host memrefs out, every process on the device;
checker = the oracle on the same bytes: leveled ops and keyswitch word for word, PBS by decryption
and phase."""
import numpy as np
import pytest

from harness import KeySet, PARAMS, expand_lut, orc, phase_distance, phase_tolerance, trivial_glwe

pytestmark = pytest.mark.gpu


@pytest.fixture()
def runtime(engine):
    from concrete_b200.runtime import Runtime

    rt = Runtime(num_gpus=1)
    yield rt
    rt.close()


def test_sdfg_keyswitch_bootstrap_chain(runtime):
    """PUT ciphertexts + table, keyswitch -> bootstrap chained on the device, GET the result;
    then GET the intermediate (kept on the device, downloaded on demand) and a second batch
    through the same graph (new generation)."""
    from concrete_b200.sdfg import Graph, TOPO_TO_X86, X86_TO_TOPO

    ks = KeySet(PARAMS["TOY1024"], seed=42)
    p = ks.p
    runtime.add_ksk(ks.ksk), runtime.add_bsk(ks.bsk)
    bits = 3
    table = (np.arange(8) ** 2 % 8).astype(np.uint64)
    lut_poly = expand_lut(table, p.N, bits)
    g = Graph(runtime, num_gpus=1)
    s_in, s_mid, s_out = g.stream(X86_TO_TOPO), g.stream(), g.stream(TOPO_TO_X86)
    s_tlu = g.stream(X86_TO_TOPO, batch=False)
    g.keyswitch(s_in, s_mid, p)
    g.bootstrap(s_mid, s_tlu, s_out, p)
    for num, seed in ((4801, 3), (77, 4)):          # pipelined chunks, then a small batch
        msgs = (np.arange(num) * 5 % 8).astype(np.uint64)
        cts = ks.encrypt_big(msgs, bits, seed=seed)
        g.put_batch(s_in, cts)
        g.put(s_tlu, lut_poly)
        out = g.get_batch(s_out, np.zeros((num, p.big_n + 1), dtype=np.uint64))
        mid = g.get_batch(s_mid, np.zeros((num, p.n + 1), dtype=np.uint64))
        ref_mid = orc.keyswitch(cts, ks.ksk, p.big_n, p.n, p.b_ks, p.l_ks)
        assert np.array_equal(mid, ref_mid)
        assert np.array_equal(ks.decrypt_big(out, bits), table[msgs])
        pick = np.arange(0, num, max(1, num // 24))
        ref_out = orc.bootstrap(np.ascontiguousarray(ref_mid[pick]), trivial_glwe(lut_poly, p.k), None,
                                ks.fbsk, p.n, p.k, p.N, p.b, p.l)
        assert phase_distance(ks, np.ascontiguousarray(out[pick]), ref_out).max() <= phase_tolerance(p)
    g.close()


def test_sdfg_leveled_then_mapped_bootstrap(runtime):
    """(a + b) * w + c, negated, then keyswitch + one table per sample: every process kind, the
    {1, num} / {1, 1} operand conventions (GPUDFG.cpp:1306-1308, 1364-1366)."""
    from concrete_b200.sdfg import Graph, TOPO_TO_X86, X86_TO_TOPO

    ks = KeySet(PARAMS["TOY512"], seed=42)
    p = ks.p
    runtime.add_ksk(ks.ksk), runtime.add_bsk(ks.bsk)
    bits, num = 3, 333
    rng = np.random.default_rng(4)
    a = rng.integers(0, 2, num).astype(np.uint64)
    b = rng.integers(0, 2, num).astype(np.uint64)
    w = rng.integers(1, 3, num).astype(np.uint64)       # per-sample cleartext multipliers
    ca, cb = ks.encrypt_big(a, bits, seed=1), ks.encrypt_big(b, bits, seed=2)
    shift = np.uint64(63 - bits)
    plain = (rng.integers(0, 2, num).astype(np.uint64) << shift)
    g = Graph(runtime, num_gpus=1)
    sa, sb, sw, sp_, sc = (g.stream(X86_TO_TOPO) for _ in range(5))
    s1, s2, s3, s4, s5, s6, s_mid, s_out = (g.stream() for _ in range(8))
    s_tlu = g.stream(X86_TO_TOPO)
    g.add(sa, sb, s1)
    g.mul_cleartext(s1, sw, s2)
    g.add_plaintext(s2, sp_, s3)
    g.negate(s3, s4)
    g.mul_cleartext(s4, sc, s5, cst=True)               # constant -1 (wrapping): back to positive
    g.add_plaintext(s5, sc, s6, cst=True)               # + (2^64 - 1): one ulp below, harmless
    g.keyswitch(s6, s_mid, p)
    g.bootstrap(s_mid, s_tlu, s_out, p, mapped=True)
    g.put_batch(sa, ca)
    g.put_batch(sb, cb)
    g.put_batch(sw, w.reshape(1, num))
    g.put_batch(sp_, plain.reshape(1, num))
    g.put_batch(sc, np.array([[2 ** 64 - 1]], dtype=np.uint64))
    tables = np.stack([np.roll(np.arange(8, dtype=np.uint64), s % 8) for s in range(num)])
    tlus = np.stack([expand_lut(t, p.N, bits) for t in tables])
    g.put_batch(s_tlu, tlus)
    out = g.get_batch(s_out, np.zeros((num, p.big_n + 1), dtype=np.uint64))
    pre = g.get_batch(s6, np.zeros((num, p.big_n + 1), dtype=np.uint64))
    # oracle side: the same wrapping arithmetic in numpy
    ref = (ca + cb) * w[:, None]
    ref[:, p.big_n] += plain
    ref = (np.uint64(0) - ref) * np.uint64(2 ** 64 - 1)
    ref[:, p.big_n] += np.uint64(2 ** 64 - 1)
    assert np.array_equal(pre, ref)
    value = ((a + b) * w + (plain >> shift)) % 8
    assert np.array_equal(ks.decrypt_big(out, bits), tables[np.arange(num), value])
    mid = g.get_batch(s_mid, np.zeros((num, p.n + 1), dtype=np.uint64))
    assert np.array_equal(mid, orc.keyswitch(ref, ks.ksk, p.big_n, p.n, p.b_ks, p.l_ks))
    g.close()


def test_sdfg_sharded_over_two_gpus(runtime):
    """SDFG_NUM_GPUS > 1: contiguous shards, one host thread + stream per device."""
    import torch
    from concrete_b200.sdfg import Graph, TOPO_TO_X86, X86_TO_TOPO

    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    runtime.lib.concrete_b200_context_set_num_gpus(runtime.ctx, 2)
    ks = KeySet(PARAMS["TOY1024"], seed=42)
    p = ks.p
    runtime.add_ksk(ks.ksk), runtime.add_bsk(ks.bsk)
    bits, num = 3, 2 * 1184 * 2 + 101
    table = (np.arange(8) ** 2 % 8).astype(np.uint64)
    g = Graph(runtime, num_gpus=2)
    s_in, s_mid, s_out = g.stream(X86_TO_TOPO), g.stream(), g.stream(TOPO_TO_X86)
    s_tlu = g.stream(X86_TO_TOPO, batch=False)
    g.keyswitch(s_in, s_mid, p)
    g.bootstrap(s_mid, s_tlu, s_out, p)
    msgs = (np.arange(num) * 3 % 8).astype(np.uint64)
    cts = ks.encrypt_big(msgs, bits)
    g.put_batch(s_in, cts)
    g.put(s_tlu, expand_lut(table, p.N, bits))
    out = g.get_batch(s_out, np.zeros((num, p.big_n + 1), dtype=np.uint64))
    assert np.array_equal(ks.decrypt_big(out, bits), table[msgs])
    mid = g.get_batch(s_mid, np.zeros((num, p.n + 1), dtype=np.uint64))
    assert np.array_equal(mid, orc.keyswitch(cts, ks.ksk, p.big_n, p.n, p.b_ks, p.l_ks))
    g.close()
