#!/usr/bin/env python
"""Opcode mix and hottest source lines of one kernel in an ncu report: ncu_opmix.py report.ncu-rep [kernel-id]"""
import csv, subprocess, sys, collections
rep = sys.argv[1]
kid = sys.argv[2] if len(sys.argv) > 2 else None
cmd = ['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'sass,cuda']
if kid is not None: cmd += ['--kernel-id', ':::%s' % kid] if not kid.isdigit() else ['--launch-skip', kid, '--launch-count', '1']
out = subprocess.run(cmd, capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
cur = None; hdr = None; lines = []; ops = collections.Counter(); tot = 0
for r in rows:
    if r and r[0] == 'File Path': cur = r[1].split('/')[-1]; continue
    if r and r[0] == 'Function Name': print('kernel:', r[1]); continue
    if r and r[0] == 'Line No': hdr = r; ie = r.index('Instructions Executed'); continue
    if not hdr or not r: continue
    off = len(hdr) - ie
    if r[0].isdigit():
        try: n = int(r[len(r) - off])
        except ValueError: continue
        try: smp = int(r[len(r) - off - 1])
        except ValueError: smp = 0
        lines.append((n, cur, r[0], ','.join(r[1:len(r) - len(hdr) + 2])[:110], smp))
    elif r[0] == '' and len(r) > 3 and r[2].startswith('0x'):
        try: n = int(r[len(r) - off])
        except ValueError: continue
        t = r[3].split()
        op = t[1] if t[0].startswith('@') else t[0]
        ops[op.split('.')[0]] += n; tot += n
print('total warp instructions', tot)
for k, v in ops.most_common(16): print('  %-10s %10d %5.1f%%' % (k, v, 100.0 * v / max(tot, 1)))
lines.sort(reverse=True)
for l in lines[:28]: print(l)
print('--- by stall samples')
tots = sum(l[4] for l in lines)
for l in sorted(lines, key=lambda l: -l[4])[:22]: print('%5.1f%%' % (100.0 * l[4] / max(tots, 1)), l[1], l[2], l[3][:100])
