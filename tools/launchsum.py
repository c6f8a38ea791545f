import csv, collections, sys
rows = list(csv.reader(l for l in open(sys.argv[1]) if not l.startswith('==')))
hdr = rows[0]; ki = hdr.index('Kernel Name'); vi = hdr.index('Metric Value'); ui = hdr.index('Metric Unit')
seq = [(r[ki].split('(')[0], float(r[vi].replace(',', '')) / (1000 if r[ui] == 'ns' else 1)) for r in rows[1:] if len(r) > vi]
agg = collections.defaultdict(list)
for k, v in seq: agg[k].append(v)
tot = sum(v for k, v in seq)
npic = max(1, len(agg.get('st_finish_kernel', agg.get('st_init_kernel', [1]))))
print('pictures %d, serialized device time %.1f us/picture' % (npic, tot / npic))
for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
    print('%-40s n/pic %4.1f  avg %6.1f us  per-picture %6.1f us (%4.1f%%)' % (k[-40:], len(v) / npic, sum(v) / len(v), sum(v) / npic, 100 * sum(v) / tot))
