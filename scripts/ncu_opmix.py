#!/usr/bin/env python
"""Opcode mix of one kernel from an .ncu-rep source page.  usage: ncu_opmix.py rep kernel_regex [top]"""
import collections, csv, io, subprocess, sys
rep, pat = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 24
raw = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--kernel-name', 'regex:' + pat, '--launch-count', '1'],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = rows[1]
i_src, i_ex, i_th, i_s = (hdr.index(x) for x in ('Source', 'Instructions Executed', 'Thread Instructions Executed', '# Samples'))
ops, samp, th, tot = collections.Counter(), collections.Counter(), collections.Counter(), 0
for r in rows[2:]:
    try:
        ex = int(r[i_ex])
    except Exception:
        continue
    toks = r[i_src].split()
    op = (toks[1] if toks[0].startswith('@') else toks[0]).split('.')[0]
    ops[op] += ex; samp[op] += int(r[i_s]); tot += ex; th[op] += int(r[i_th])
print('total warp inst', tot, 'samples', sum(samp.values()))
for op, c in ops.most_common(top):
    print(f'{op:10s} {c:12d} {100*c/tot:5.1f}%  samples {samp[op]:7d}  avg thr {th[op]/max(c,1):.1f}')
