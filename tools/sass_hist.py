"""Summarise an `ncu --page source --csv --print-source sass` dump: opcode histogram weighted by executed count, stall totals."""
import collections
import csv
import re
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr_idx = [i for i, r in enumerate(rows) if r and r[0] == 'Address']
h = rows[hdr_idx[0]]
end = hdr_idx[1] - 1 if len(hdr_idx) > 1 else len(rows)
body = [r for r in rows[hdr_idx[0] + 1:end] if len(r) == len(h)]
ci, ti, si = h.index('Instructions Executed'), h.index('Thread Instructions Executed'), h.index('# Samples')
tot_i = sum(int(r[ci]) for r in body)
tot_t = sum(int(r[ti]) for r in body)
print("kernel:", rows[0][1][:100])
print("warp instrs", tot_i, "thread instrs", tot_t, "avg active threads", round(tot_t / tot_i, 2), "samples", sum(int(r[si]) for r in body), "sass lines", len(body))
op = collections.Counter()
for r in body:
    m = re.match(r'\s*(@!?U?P\d+\s+)?([A-Z0-9_.]+)', r[1])
    op[m.group(2).split('.')[0] if m else '?'] += int(r[ci])
for o, c in op.most_common(28):
    print(f"  {o:10s} {c:10d} {100 * c / tot_i:5.1f}%")
for name in ['stall_long_sb', 'stall_short_sb', 'stall_wait', 'stall_math', 'stall_not_selected', 'stall_selected', 'stall_lg', 'stall_branch_resolving',
             'stall_no_inst', 'stall_dispatch', 'stall_mio', 'stall_tex']:
    i = h.index(name)
    print(f"  {name:24s} {sum(int(r[i]) for r in body)}")
