# DRAM bytes of the search kernels per picture (ncu, cold caches): writes profiles/traffic.json
ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -s 2000 -c 640 --csv --log-file gpurun_out/traffic.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_t.log 2>&1
python tools/trafficsum.py gpurun_out/traffic.csv gpurun_out/traffic.json
