"""Top stall hotspots of an `ncu --page source --csv` dump.  usage: ncu_hotspots.py src.csv [N]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 15
hdr = rows[1]; ci = {h: i for i, h in enumerate(hdr)}
tot = 0; data = []
for r in rows[2:]:
    try: v = float(r[ci['# Samples']])
    except Exception: continue
    tot += v
    stalls = {k: float(r[ci[k]] or 0) for k in hdr if k.startswith('stall_') and 'Not Issued' not in k and r[ci[k]] not in ('', '0')}
    top = sorted(stalls.items(), key=lambda kv: -kv[1])[:2]
    data.append((v, r[ci['Source']][:100], top))
data.sort(key=lambda t: -t[0])
print("total samples", tot)
for v, s, top in data[:n]:
    print(f"{v:7.0f} {100*v/tot:5.1f}%  {s}   {top}")
