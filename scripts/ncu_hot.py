#!/usr/bin/env python
"""Hottest SASS lines of one kernel by a stall column.  usage: ncu_hot.py rep kernel_regex [column=# Samples] [top=25]"""
import csv, io, subprocess, sys
rep, pat = sys.argv[1], sys.argv[2]
col = sys.argv[3] if len(sys.argv) > 3 else '# Samples'
top = int(sys.argv[4]) if len(sys.argv) > 4 else 25
raw = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--kernel-name', 'regex:' + pat, '--launch-count', '1'],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = rows[1]
ic, isrc, iex, ia = hdr.index(col), hdr.index('Source'), hdr.index('Instructions Executed'), hdr.index('Address')
body = [r for r in rows[2:] if len(r) > ic]
def val(r):
    try: return int(r[ic])
    except Exception: return 0
order = sorted(range(len(body)), key=lambda i: -val(body[i]))[:top]
for i in sorted(order):
    r = body[i]
    print(f'{i:5d} {val(r):6d} ex={r[iex]:>9s}  {r[isrc][:110]}')
