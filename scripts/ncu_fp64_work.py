#!/usr/bin/env python
"""FP64 operations a kernel executes, from the source page of an .ncu-rep (thread-level: only active threads count).
DFMA = 2 flop, DMUL / DADD = 1; DSETP, conversions and MUFU are pipe time but not arithmetic and are listed apart.
usage: ncu_fp64_work.py rep kernel_regex units_per_launch   (units: face points, or DOF of the column)"""
import collections, csv, io, subprocess, sys
rep, pat, units = sys.argv[1], sys.argv[2], float(sys.argv[3])
raw = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--kernel-name', 'regex:' + pat, '--launch-count', '1'],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = rows[1]
i_src, i_ex, i_th = (hdr.index(x) for x in ('Source', 'Instructions Executed', 'Thread Instructions Executed'))
warp, thread = collections.Counter(), collections.Counter()
for r in rows[2:]:
    try:
        ex, th = int(r[i_ex]), int(r[i_th])
    except Exception:
        continue
    toks = r[i_src].split()
    op = (toks[1] if toks[0].startswith('@') else toks[0]).split('.')[0]
    warp[op] += ex; thread[op] += th
flop = 2 * thread['DFMA'] + thread['DMUL'] + thread['DADD']
pipe = warp['DFMA'] + warp['DMUL'] + warp['DADD'] + warp['DSETP']
print({'kernel': pat, 'flop_per_unit': flop / units, 'flop_per_launch': flop, 'warp_inst_total': sum(warp.values()),
       'warp_inst_fp64_pipe': pipe, 'thread_DFMA': thread['DFMA'], 'thread_DMUL': thread['DMUL'], 'thread_DADD': thread['DADD']})
