"""Phase-distance statistics of the PBS kernel families against the oracle on one input set
(development helper; run on the GPU box): python tools/p8_phase_stats.py [PARAMS] [n] [num]"""
import sys
sys.path.insert(0, "tests"); sys.path.insert(0, ".")
import numpy as np
from harness import KeySet, PARAMS, expand_lut, orc, trivial_glwe, phase_distance, phase_tolerance
from concrete_b200 import Engine, as_u64

name = sys.argv[1] if len(sys.argv) > 1 else "P8"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 3
num = int(sys.argv[3]) if len(sys.argv) > 3 else 177
ks = KeySet(PARAMS[name], seed=42, n_override=n)
p = ks.p
bits = 4
rng = np.random.default_rng(13)
msgs = rng.integers(0, 1 << bits, num).astype(np.uint64)
table = (np.arange(1 << bits) ** 2 % (1 << bits)).astype(np.uint64)
lut = trivial_glwe(expand_lut(table, p.N, bits), p.k)
mid = ks.encrypt_small(msgs, bits, seed=13)
ref = orc.bootstrap(mid, lut, None, ks.fbsk, p.n, p.k, p.N, p.b, p.l)
eng = Engine(0)
outs = {}
for plan in (0, 1, 3, 0):
    eng.lib.concrete_b200_set_pbs_plan(plan)
    d_fbsk = eng.convert_bsk(ks.bsk, p.n, p.k, p.l, p.N)
    got = as_u64(eng.bootstrap(eng.to_device(mid), eng.to_device(lut), None, d_fbsk, p.n, p.k, p.N, p.b, p.l))
    d = phase_distance(ks, got, ref)
    ok = np.array_equal(ks.decrypt_big(got, bits), table[msgs])
    print(f"plan {plan}: decrypt ok {ok}  log2 max {np.log2(d.max()+1):.2f}  log2 rms {np.log2(np.sqrt((d**2).mean())+1):.2f}  "
          f"tolerance 2^{np.log2(phase_tolerance(p)):.2f}  argmax {d.argmax()}", flush=True)
    if plan in outs:
        print("   deterministic (bit-identical to the first run of this plan):", np.array_equal(outs[plan], got))
    outs[plan] = got
for a, b in ((0, 1), (0, 3), (1, 3)):
    d = phase_distance(ks, outs[a], outs[b])
    print(f"plan {a} vs plan {b}: log2 max {np.log2(d.max()+1):.2f} log2 rms {np.log2(np.sqrt((d**2).mean())+1):.2f}")
eng.lib.concrete_b200_set_pbs_plan(0)
