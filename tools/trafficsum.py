import csv, collections, json, sys
rows = list(csv.reader(l for l in open(sys.argv[1]) if not l.startswith('==')))
hdr = rows[0]; ki = hdr.index('Kernel Name'); vi = hdr.index('Metric Value'); ui = hdr.index('Metric Unit'); ni = hdr.index('Metric Name')
mult = {'byte': 1, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}
per = collections.defaultdict(float); cnt = collections.Counter()
for r in rows[1:]:
    if len(r) <= vi: continue
    k = r[ki].split('(')[0]
    per[k] += float(r[vi].replace(',', '')) * mult.get(r[ui], 1)
    if r[ni] == 'dram__bytes_read.sum': cnt[k] += 1
npic = max(1, cnt.get('st_finish_kernel', cnt.get('st_init_kernel', 1)))  # one st_init_kernel per picture search
search = sum(v for k, v in per.items() if k.split('<')[0].split()[-1].startswith('st_'))
out = {"search_dram_bytes_per_picture": search / npic, "pictures": npic,
       "per_kernel_bytes_per_picture": {k: v / npic for k, v in sorted(per.items(), key=lambda kv: -kv[1])},
       "source": "ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum over %d launches of bench.py (cold caches per launch)" % sum(cnt.values())}
json.dump(out, open(sys.argv[2], 'w'), indent=1)
print(json.dumps(out, indent=1))
