# A/B of library builds: every variants/*.so through the default bench, REPS times interleaved
for r in $(seq ${REPS:-1}); do for f in variants/*.so; do echo "== $f"; SVTME_LIB=$PWD/$f python bench.py --steps ${STEPS:-3} --warmup 3 --no-cpu-baseline ${BENCH_ARGS} 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('value %.0f e2e %.0f frac %.3f launch_ms %.3f'%(d['value'],d['e2e']['value'],d['roofline']['serialised']['frac'],d['roofline']['serialised']['avg_launch_ms']))
    else: print(l.rstrip()[-300:])
"; done; done
