# A/B of environment switches: bash tools/abenv.sh "VAR=a" "VAR=b" ...  (REPS interleaved repetitions of the default bench)
for r in $(seq ${REPS:-2}); do for e in "$@"; do echo "== $e"; env $e python bench.py --steps ${STEPS:-3} --warmup 3 --no-cpu-baseline ${BENCH_ARGS} 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('value %.0f e2e %.0f frac %.3f launch_ms %.3f'%(d['value'],d['e2e']['value'],d['roofline']['serialised']['frac'],d['roofline']['serialised']['avg_launch_ms']))
    else: print(l.rstrip()[-300:])
"; done; done
