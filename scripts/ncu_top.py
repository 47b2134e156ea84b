"""Summarise an `ncu --page source --csv` dump: stall-reason totals and the hottest SASS lines."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 30
hi = [i for i, r in enumerate(rows) if r and r[0] == 'Address'][0]
hdr = rows[hi]
data = [r for r in rows[hi + 1:] if len(r) == len(hdr) and r[0] != 'Address']
ix = {h: i for i, h in enumerate(hdr)}
def num(r, k):
    try: return int(float(r[ix[k]] or 0))
    except Exception: return 0
stalls = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
tot = sum(num(r, '# Samples') for r in data)
print("total samples", tot, "instructions", len(data))
agg = {s: sum(num(r, s) for r in data) for s in stalls}
print([(k, v, f"{100*v/max(tot,1):.1f}%") for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]])
for r in sorted(data, key=lambda r: -num(r, '# Samples'))[:n]:
    st = sorted(((s, num(r, s)) for s in stalls), key=lambda kv: -kv[1])[:2]
    print(str(num(r, '# Samples')).rjust(7), r[ix['Source']][:100].ljust(100), st)
