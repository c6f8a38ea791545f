#!/usr/bin/env python
"""Stall samples and executed instructions per source file (and the hottest lines of one file) of one launch in an ncu
report: ncu_byfile.py report.ncu-rep launch-index [file]"""
import csv, subprocess, sys, collections
rep = sys.argv[1]; skip = sys.argv[2]
cmd = ['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'sass,cuda', '--launch-skip', skip, '--launch-count', '1']
out = subprocess.run(cmd, capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
cur = None; hdr = None
byfile = collections.Counter(); insts = collections.Counter(); lines = []
for r in rows:
    if r and r[0] == 'File Path': cur = r[1].split('/')[-1]; continue
    if r and r[0] == 'Line No': hdr = r; ie = r.index('Instructions Executed'); continue
    if not hdr or not r or not r[0].isdigit(): continue
    off = len(hdr) - ie
    try: n = int(r[len(r) - off])
    except ValueError: continue
    try: smp = int(r[len(r) - off - 1])
    except ValueError: smp = 0
    byfile[cur] += smp; insts[cur] += n
    lines.append((smp, n, cur, r[0], ','.join(r[1:len(r) - len(hdr) + 2])[:100]))
tot = max(1, sum(byfile.values())); ti = max(1, sum(insts.values()))
for k, v in byfile.most_common(): print('%-28s samples %6.1f%%  inst %6.1f%%' % (k, 100 * v / tot, 100 * insts[k] / ti))
if len(sys.argv) > 3:
    for l in sorted([l for l in lines if l[2] == sys.argv[3]], reverse=True)[:25]: print('%5.2f%%' % (100 * l[0] / tot), l[1:])
