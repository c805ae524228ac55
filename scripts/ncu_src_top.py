#!/usr/bin/env python3
"""Summarise `ncu --page source --csv` output: top SASS instructions by stall samples (first kernel only)."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
topn = int(sys.argv[2]) if len(sys.argv) > 2 else 50
# first kernel block
start = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
end = next((i for i in range(start + 1, len(rows)) if rows[i] and rows[i][0] in ("Kernel Name", "Address")), len(rows))
hdr = rows[start]
data = [r for r in rows[start + 1:end] if len(r) == len(hdr)]
ci = {h: i for i, h in enumerate(hdr)}
tot = sum(int(r[ci['# Samples']]) for r in data)
print('total samples', tot, 'instructions', len(data))
stalls = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
agg = {s: sum(int(r[ci[s]]) for r in data) for s in stalls}
print('stall totals', sorted(agg.items(), key=lambda kv: -kv[1])[:8])
idx = sorted(range(len(data)), key=lambda i: -int(data[i][ci['# Samples']]))[:topn]
for i in sorted(idx):
    r = data[i]
    st = {s: int(r[ci[s]]) for s in stalls if int(r[ci[s]]) > 0}
    top = sorted(st.items(), key=lambda kv: -kv[1])[:3]
    print(i, r[ci['Source']].strip()[:64], r[ci['# Samples']], r[ci['Instructions Executed']], top)
