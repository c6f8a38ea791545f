# searches in flight (streams) against the picture rate: SWEEP="4 8" BENCH_ARGS="--config 1" bash tools/inflight_sweep.sh
for i in ${SWEEP:-4 6 8 12 16 24}; do echo "== inflight $i"; python bench.py --steps 3 --warmup 3 --no-cpu-baseline --inflight $i ${BENCH_ARGS} 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('value %.0f e2e %.0f'%(d['value'],d['e2e']['value']))
"; done
