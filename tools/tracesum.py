"""Aggregate SVTME_TRACE=1 lines (diagnostics build) per kernel label: python tools/tracesum.py < stderr.log"""
import sys, re, collections
tot = collections.defaultdict(float); cnt = collections.Counter(); npic = 0; grand = 0.0
for l in sys.stdin:
    if not l.startswith('[svtme trace]'): continue
    npic += 1
    body = l.split(':', 1)[1].split('|')[0]
    for name, us in re.findall(r'(\w+) ([\d.]+)', body):
        tot[name] += float(us); cnt[name] += 1; grand += float(us)
print('pictures %d, %.1f us/picture' % (npic, grand / max(npic, 1)))
for k, v in sorted(tot.items(), key=lambda kv: -kv[1]):
    print('%-8s n/pic %4.1f  avg %6.1f us  per-picture %6.1f us' % (k, cnt[k] / npic, v / cnt[k], v / npic))
