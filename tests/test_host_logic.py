"""Host-side logic: parameter sets and roofline counts, shard rule, and the N>1 path over gloo
(world_size 2, CPU): keys broadcast once from rank 0, contiguous ciphertext shards, no data-path
collective, results gathered — with the oracle standing in for the device kernels."""
import os
import sys

import numpy as np
import pytest

from concrete_b200.params import PARAMS, security_sigma
from concrete_b200.sharding import shard_counts, shard_range


def test_parameter_sets_match_survey_table():
    p = PARAMS["P2048"]
    assert (p.n, p.k, p.N, p.l, p.b, p.l_ks, p.b_ks) == (742, 1, 2048, 1, 23, 5, 3)
    assert round(p.bsk_bytes / 1e6, 1) == 48.6 and round(p.ksk_bytes / 1e6, 1) == 60.9
    p8 = PARAMS["P8"]
    assert round(p8.bsk_bytes / 1e6) == 1055 and round(p8.ksk_bytes / 1e6) == 660
    assert round(PARAMS["P4"].bsk_bytes / 1e6, 1) == 59.1


def test_roofline_counts():
    p = PARAMS["P2048"]
    # 262 144 flop per CMUX iteration, 1.945e8 per PBS
    assert p.pbs_flops() / p.n == 262144
    assert abs(p.pbs_flops() - 1.945e8) / 1.945e8 < 1e-3
    assert abs(PARAMS["P8"].pbs_flops() - 4.04e9) / 4.04e9 < 5e-3
    assert abs(p.pbs_bytes(16384) - 414e6) / 414e6 < 5e-3
    assert p.ks_macs() == 2048 * 5 * 743


def test_security_sigma_curve():
    assert security_sigma(742) == pytest.approx(2 ** (-0.025696778711484593 * 742 + 2.675931372549016))
    assert security_sigma(10 ** 6) == 2.0 ** -62


def test_shard_rule():
    assert shard_counts(10, 4) == [3, 3, 2, 2]
    assert shard_counts(3, 8) == [1, 1, 1, 0, 0, 0, 0, 0]
    assert shard_counts(0, 2) == [0, 0]
    for total in (0, 1, 7, 4096, 16385):
        for parts in (1, 2, 3, 8):
            counts = shard_counts(total, parts)
            assert sum(counts) == total and max(counts) - min(counts) <= 1
            begin = 0
            for g in range(parts):
                assert shard_range(total, parts, g) == (begin, counts[g])
                begin += counts[g]
    with pytest.raises(ValueError):
        shard_counts(4, 0)
    with pytest.raises(ValueError):
        shard_range(4, 2, 2)


def _gloo_worker(rank, world, port, tmpdir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank),
                      WORLD_SIZE=str(world))
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    import torch
    import torch.distributed as dist

    from concrete_b200.sharding import broadcast_keys, shard_range
    from harness import KeySet, PARAMS, expand_lut, orc, trivial_glwe

    dist.init_process_group("gloo", rank=rank, world_size=world)
    p = PARAMS["TOY1024"]
    bits, total = 3, 13  # ragged: 7 + 6
    ks = KeySet(p, seed=42, fourier_oracle=(rank == 0))
    ksk = torch.from_numpy(ks.ksk.view(np.int64).copy()) if rank == 0 else torch.zeros(ks.ksk.size, dtype=torch.int64)
    fbsk = (torch.from_numpy(ks.fbsk.view(np.float64).copy()) if rank == 0
            else torch.zeros(ks.bsk.size, dtype=torch.float64))
    broadcast_keys([ksk, fbsk], src=0)  # the one collective: keys, once
    ksk_np = ksk.numpy().view(np.uint64)
    fbsk_np = fbsk.numpy().view(np.complex128)
    msgs = (np.arange(total) % 8).astype(np.uint64)
    cts = ks.encrypt_big(msgs, bits, seed=5)  # same seed on every rank: identical global batch
    begin, cnt = shard_range(total, world, rank)
    table = (np.arange(8) ** 2 % 8).astype(np.uint64)
    lut = trivial_glwe(expand_lut(table, p.N, bits), p.k)
    mid = orc.keyswitch(np.ascontiguousarray(cts[begin:begin + cnt]), ksk_np, p.big_n, p.n, p.b_ks, p.l_ks)
    out = orc.bootstrap(mid, lut, None, fbsk_np, p.n, p.k, p.N, p.b, p.l)
    np.save(os.path.join(tmpdir, f"out{rank}.npy"), out)
    dist.barrier()
    if rank == 0:
        full = np.concatenate([np.load(os.path.join(tmpdir, f"out{r}.npy")) for r in range(world)])
        ref_mid = orc.keyswitch(cts, ks.ksk, p.big_n, p.n, p.b_ks, p.l_ks)
        ref = orc.bootstrap(ref_mid, lut, None, ks.fbsk, p.n, p.k, p.N, p.b, p.l)
        assert np.array_equal(full, ref)
        assert np.array_equal(ks.decrypt_big(full, bits), table[msgs])
    dist.destroy_process_group()


def test_sharded_path_world_size_2_gloo(tmp_path):
    import socket

    import torch.multiprocessing as mp

    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mp.spawn(_gloo_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
