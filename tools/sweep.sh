cd "$(dirname "$0")/.." && true
run() { env "$@" timeout 300 python bench.py --steps 2 --warmup 1 --samples ${SAMPLES:-7104} --no-cpu-baseline --no-e2e --no-ref-gpu 2>&1 | python -c "
import sys,json
tag=' '.join(sys.argv[1:])
ok=False
for l in sys.stdin:
    if l.startswith('{'):
        j=json.loads(l); ok=True; print(tag, 'pbs_ms', round(j['roofline']['launch_ms'],2), 'frac', round(j['roofline']['frac'],4), 'ok', j['verified_decryptions'])
    elif 'rror' in l or 'panic' in l: print(tag, l.strip()[:200])
if not ok: print(tag,'NO RESULT')
" "$@"; }
for v in "$@"; do run $v; done
