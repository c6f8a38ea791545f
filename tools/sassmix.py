#!/usr/bin/env python
"""Pipe mix of the SAD loop of a kernel, from the SASS of the built library:
   python tools/sassmix.py <kernel-name-substring> <number of VABSDIFF4 in the loop body>
ALU pipe (half rate, shared with the SADs): IADD3 LOP3 SHF PRMT ISETP SEL VIMNMX VIADD LEA IABS ...; FMA pipe: IMAD*, FFMA...;
LSU/MIO: LDS STS LDG LDGSTS SHFL ATOMS."""
import collections, re, subprocess, sys
lib = 'svt-av1-mirror_b200/svtme/libsvtme_cuda.so'
out = subprocess.run(['cuobjdump', '-sass', lib], capture_output=True, text=True).stdout
funcs, cur = {}, None
for l in out.splitlines():
    m = re.match(r'\s*Function : (\S+)', l)
    if m: cur = m.group(1); funcs[cur] = []; continue
    m = re.match(r'\s*/\*([0-9a-f]{4,5})\*/\s+(.*?);', l)
    if m and cur: funcs[cur].append(m.group(2))
ALU = ('IADD3', 'LOP3', 'SHF', 'PRMT', 'ISETP', 'SEL', 'VIMNMX', 'VIADD', 'LEA', 'IABS', 'VABSDIFF4', 'VABSDIFF', 'PLOP3', 'IMNMX', 'POPC', 'FLO', 'BREV', 'UIADD3', 'ULOP3', 'USHF', 'ULEA', 'R2P', 'P2R', 'CS2R', 'MOV', 'FMNMX', 'FSEL', 'FSETP')
FMA = ('IMAD', 'FFMA', 'FMUL', 'FADD', 'HFMA2', 'UIMAD')
LSU = ('LDS', 'STS', 'LDG', 'STG', 'LDGSTS', 'SHFL', 'ATOMS', 'ATOM', 'RED', 'LDC', 'LDSM', 'ATOMG')
for name, ops in funcs.items():
    if sys.argv[1] not in name: continue
    n = int(sys.argv[2])
    idx = [i for i, o in enumerate(ops) if 'VABSDIFF4' in o]
    if len(idx) < n: print(name, 'has only', len(idx), 'VABSDIFF4'); continue
    span, s = min((idx[k + n - 1] - idx[k], k) for k in range(len(idx) - n + 1))
    body = ops[idx[s]:idx[s + n - 1] + 1]
    c = collections.Counter()
    for o in body:
        t = o.split(); op = (t[1] if t[0].startswith('@') else t[0]).split('.')[0]
        c[op] += 1
    alu = sum(v for k, v in c.items() if k in ALU); fma = sum(v for k, v in c.items() if k in FMA); lsu = sum(v for k, v in c.items() if k in LSU)
    print('%s\n  densest span with %d SADs: %d instrs; ALU pipe %d (SAD share %.0f%%), FMA pipe %d, LSU %d, other %d' % (
        name[:90], n, len(body), alu, 100.0 * n / alu, fma, lsu, len(body) - alu - fma - lsu))
    print('  ', dict(c.most_common()))
