#!/usr/bin/env python
"""Stall-reason totals of one kernel from the source page of an .ncu-rep.  usage: ncu_stalls.py rep kernel_regex"""
import collections, csv, io, subprocess, sys
rep, pat = sys.argv[1], sys.argv[2]
raw = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--kernel-name', 'regex:' + pat, '--launch-count', '1'],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = rows[1]
cols = [i for i, h in enumerate(hdr) if h.startswith('stall_') and 'Not Issued' not in h]
tot = collections.Counter()
for r in rows[2:]:
    for i in cols:
        try: tot[hdr[i]] += int(r[i])
        except Exception: pass
s = sum(tot.values())
print('samples', s)
for k, v in tot.most_common(): print(f'{k:24s} {v:8d} {100*v/max(s,1):5.1f}%')
