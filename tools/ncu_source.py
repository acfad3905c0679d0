"""Stall samples per code segment from an .ncu-rep source page: python tools/ncu_source.py rep [seg]"""
import csv, subprocess, sys
rep = sys.argv[1]; seg = int(sys.argv[2]) if len(sys.argv) > 2 else 400
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = rows[1]; data = rows[2:]
ix = {h: i for i, h in enumerate(hdr)}
def f(r, k):
    try: return float(r[ix[k]])
    except Exception: return 0.0
tot = sum(f(r, '# Samples') for r in data)
print('total samples', tot, 'instructions', len(data))
stalls = ['stall_long_sb', 'stall_wait', 'stall_no_inst', 'stall_math', 'stall_not_selected', 'stall_short_sb',
          'stall_selected', 'stall_dispatch', 'stall_mio', 'stall_barrier', 'stall_lg', 'stall_sleep', 'stall_branch_resolving']
for s0 in range(0, len(data), seg):
    chunk = data[s0:s0 + seg]
    smp = sum(f(r, '# Samples') for r in chunk)
    ex = sum(f(r, 'Instructions Executed') for r in chunk)
    ops = {}
    for r in chunk:
        t = r[ix['Source']].split()
        if not t: continue
        op = t[1] if t[0].startswith('@') and len(t) > 1 else t[0]
        ops[op.split('.')[0]] = ops.get(op.split('.')[0], 0) + 1
    top = sorted(ops.items(), key=lambda x: -x[1])[:4]
    st = {k: sum(f(r, k) for r in chunk) for k in stalls}
    sts = ' '.join(f"{k[6:]}={v / max(smp, 1) * 100:.0f}%" for k, v in sorted(st.items(), key=lambda x: -x[1])[:4])
    print(f"{s0:6d} samples {smp / tot * 100:5.1f}% exec {ex / 1e6:8.1f}M  {top}  {sts}")
