for pl in ${PIPES:-0 1}; do echo "== pipeline $pl"; SVTME_PIPELINE=$pl python bench.py --steps 3 --warmup 3 --no-cpu-baseline ${BENCH_ARGS} 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('value %.0f e2e %.0f frac %.3f launch_ms %.3f launches %d'%(d['value'],d['e2e']['value'],d['roofline']['serialised']['frac'],d['roofline']['serialised']['avg_launch_ms'],d['gpu_launches']))
    else: print(l.rstrip()[-300:])
"; done
