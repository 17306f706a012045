#!/usr/bin/env python
"""Per-source-line share of warp-stall samples of one kernel, from `ncu -i X.ncu-rep --page source --csv
--print-source cuda,sass --kernel-name regex:K` written to a file.  Usage: ncu_lines.py dump.csv [top]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
cur_file, cur_line, cur_src = None, None, None
agg = {}
col = None
for r in rows:
    if not r: continue
    if r[0] == 'File Path': cur_file = r[1].split('/')[-1]; continue
    if r[0] == 'Line No': col = r.index('Warp Stall Sampling (All Samples)'); continue
    if r[0] == 'Function Name' or col is None: continue
    if r[0] != '': cur_line, cur_src = r[0], r[1]
    try: n = int(r[col] or 0)
    except ValueError: continue
    if r[0] == '' or True:
        k = (cur_file, cur_line)
        if r[0] == '':   # sass row
            a = agg.setdefault(k, [0, cur_src]); a[0] += n
tot = sum(v[0] for v in agg.values()) or 1
for (f, l), (n, s) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print("%5.1f%%  %s:%s  %s" % (100.0 * n / tot, f, l, s.strip()[:120]))
print("total samples", tot)
