"""Quick manual GPU check (development aid): python tools/gpu_check.py [param names]."""
import sys, time
import numpy as np, torch
import os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
from harness import *
from concrete_b200 import Engine, as_i64, as_u64

def main():
    names = sys.argv[1:] or ["TOY512", "TOY256K2", "TOY1024"]
    e = Engine(0)
    for name in names:
        n_over = None
        if ":" in name:
            name, n_over = name.split(":"); n_over = int(n_over)
        p = PARAMS[name]
        t = time.time(); ks = KeySet(p, n_override=n_over); p = ks.p
        print(f"[{p.name}] keygen {time.time()-t:.2f}s", flush=True)
        bits = min(p.msg_bits, 4); num = 37
        rng = np.random.default_rng(3)
        msgs = rng.integers(0, 1 << bits, num).astype(np.uint64)
        cts = ks.encrypt_big(msgs, bits)
        # keyswitch
        ref_ks = orc.keyswitch(cts, ks.ksk, p.big_n, p.n, p.b_ks, p.l_ks)
        d_ksk = e.upload_ksk(ks.ksk); d_cts = e.to_device(cts)
        got_ks = as_u64(e.keyswitch(d_cts, d_ksk, p.big_n, p.n, p.b_ks, p.l_ks))
        print("  KS bit-exact:", np.array_equal(got_ks, ref_ks), "mismatch words:", int((got_ks != ref_ks).sum()), flush=True)
        # fft mul
        a = torch.rand((5, p.N), dtype=torch.float64, device="cuda") * 2 - 1
        b = torch.rand((5, p.N), dtype=torch.float64, device="cuda") * 2 - 1
        prod = e.fourier_polynomial_mul(a, b).cpu().numpy()
        an, bn = a.cpu().numpy(), b.cpu().numpy()
        ref = np.zeros_like(an)
        for q in range(5):
            full = np.convolve(an[q], bn[q])
            ref[q] = full[:p.N]; ref[q][:p.N-1] -= full[p.N:]
        print("  FFT mul max err:", np.abs(prod - ref).max(), flush=True)
        # PBS
        table = (np.arange(1 << bits) ** 2 % (1 << bits)).astype(np.uint64)
        lut = trivial_glwe(expand_lut(table, p.N, bits), p.k)
        t = time.time(); ref_pbs = orc.bootstrap(ref_ks, lut, None, ks.fbsk, p.n, p.k, p.N, p.b, p.l); t_cpu = time.time() - t
        d_fbsk = e.convert_bsk(ks.bsk, p.n, p.k, p.l, p.N)
        d_lut = e.to_device(lut)
        got = e.bootstrap(e.to_device(ref_ks), d_lut, None, d_fbsk, p.n, p.k, p.N, p.b, p.l)
        torch.cuda.synchronize()
        got = as_u64(got)
        dist = torus_distance(got, ref_pbs)
        lg = np.log2(dist + 1)
        print("  dist log2 histogram:", np.histogram(lg, bins=[0, 20, 30, 40, 50, 55, 60, 62, 64])[0])
        bad = np.argwhere(lg > 60)
        print("  bad positions (sample, word):", bad[:10].tolist(), "count", len(bad))
        for (si, wi) in bad[:4]:
            print(f"    got={int(got[si, wi]):016x} ref={int(ref_pbs[si, wi]):016x}")
        dec = ks.decrypt_big(got, bits); dec_ref = ks.decrypt_big(ref_pbs, bits)
        print(f"  PBS decrypt ok: {np.array_equal(dec, table[msgs])} (oracle ok: {np.array_equal(dec_ref, table[msgs])}) "
              f"max|d|=2^{np.log2(dist.max()+1):.1f} sigma_fft=2^{np.log2(sigma_fft_ulp(p)):.1f} cpu {t_cpu:.2f}s", flush=True)
    e.close()

if __name__ == "__main__":
    main()
