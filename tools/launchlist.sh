# per-kernel serialized times of one timed step (ncu launch list, cold caches): bash tools/launchlist.sh [out.csv]
OUT=${1:-gpurun_out/launches.csv}
ncu --metrics gpu__time_duration.sum --clock-control none -s 2000 -c 640 --csv --log-file $OUT python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_l.log 2>&1
python tools/launchsum.py $OUT
