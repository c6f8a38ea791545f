# per-step device times of the staged pipeline for a few pictures (warm): last 34 trace lines of a short bench run
SVTME_TRACE=1 python bench.py --steps 1 --warmup 3 --no-cpu-baseline --inflight 1 2>&1 | grep "svtme trace" | tail -${1:-12}
