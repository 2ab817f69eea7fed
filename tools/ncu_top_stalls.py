"""Top stalled instructions of one kernel from an ncu report's source page.

    ncu -i rep.ncu-rep --page source --csv --kernel-name regex:NAME > src.csv ; python tools/ncu_top_stalls.py src.csv [N]
"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 40
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
data = []
for r in rows[2:]:
    if len(r) != len(hdr):
        if data:
            break                # next kernel instance: keep the first one only
        continue
    if r[ix['# Samples']].isdigit():
        data.append(r)
tot = sum(int(r[ix['# Samples']] or 0) for r in data)
print(rows[0][1][:100])
print("total samples", tot, "instructions", len(data))
stalls = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
agg = {h: sum(int(r[ix[h]] or 0) for r in data) for h in stalls}
print("by reason:", sorted(agg.items(), key=lambda kv: -kv[1])[:8])
for r in sorted(data, key=lambda r: -int(r[ix['# Samples']] or 0))[:n]:
    s = {h: int(r[ix[h]] or 0) for h in stalls}
    best = sorted(s.items(), key=lambda kv: -kv[1])[:2]
    print(r[ix['# Samples']].rjust(6), r[ix['Instructions Executed']].rjust(9), r[ix['Source']][:100].ljust(100), best)
