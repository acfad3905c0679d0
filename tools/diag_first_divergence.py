"""Diagnostic: find the first CMUX iteration where GPU and oracle PBS diverge."""
import sys
import numpy as np, torch
import os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
from harness import *
from concrete_b200 import Engine, as_u64

def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "TOY1024"
    e = Engine(0)
    p = PARAMS[name]; ks = KeySet(p)
    bits = 3; num = 24
    msgs = (np.arange(num) % 8).astype(np.uint64)
    cts = ks.encrypt_big(msgs, bits)
    mid = orc.keyswitch(cts, ks.ksk, p.big_n, p.n, p.b_ks, p.l_ks)
    table = (np.arange(8) ** 2 % 8).astype(np.uint64)
    lut = trivial_glwe(expand_lut(table, p.N, bits), p.k)
    d_fbsk = e.convert_bsk(ks.bsk, p.n, p.k, p.l, p.N)
    d_lut = e.to_device(lut)
    logN = int(np.log2(p.N))
    ahat = ((mid + (np.uint64(1) << np.uint64(63 - (logN + 1)))) >> np.uint64(64 - (logN + 1)))
    def run(nn):
        sub = np.ascontiguousarray(np.concatenate([mid[:, :nn], mid[:, p.n:]], axis=1))
        ref = orc.bootstrap(sub, lut, None, ks.fbsk, nn, p.k, p.N, p.b, p.l)
        got = as_u64(e.bootstrap(e.to_device(sub), d_lut, None, d_fbsk, nn, p.k, p.N, p.b, p.l))
        return torus_distance(got, ref).max(axis=1)
    full = run(p.n)
    bad = np.where(full > 2.0 ** 50)[0]
    print("bad samples:", bad.tolist(), "log2 max dist per sample:", np.round(np.log2(full + 1), 1).tolist())
    prev = run(0)
    print("n=0 dist", np.log2(prev.max() + 1))
    first = {}
    for nn in range(1, p.n + 1):
        d = run(nn)
        for s in np.where(d > 2.0 ** 50)[0]:
            if s not in first:
                first[s] = nn - 1
                print(f"sample {s}: first divergence at iteration i={nn-1}: a_hat={int(ahat[s, nn-1])} lwe={int(mid[s, nn-1]):016x} s_i={int(ks.sk_small[nn-1])} b_hat={int(ahat[s, p.n])}")
    e.close()

main()
