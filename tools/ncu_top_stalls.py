"""Summarise an `ncu --page source --csv` dump: top SASS instructions by stall samples."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]
col = {h: i for i, h in enumerate(hdr)}
data = [r for r in rows[2:] if len(r) == len(hdr) and r[hdr.index("# Samples")].isdigit()]
stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
tot = sum(int(r[col["# Samples"]] or 0) for r in data)
print("total samples", tot)
agg = {}
for h in stall_cols:
    agg[h] = sum(int(r[col[h]] or 0) for r in data)
print({k: v for k, v in sorted(agg.items(), key=lambda kv: -kv[1]) if v})
top = sorted(data, key=lambda r: -int(r[col["# Samples"]] or 0))[: int(sys.argv[2]) if len(sys.argv) > 2 else 25]
for r in top:
    st = {h[6:]: int(r[col[h]] or 0) for h in stall_cols if int(r[col[h]] or 0)}
    print(r[col["# Samples"]].rjust(6), r[col["Instructions Executed"]].rjust(8), r[col["Source"]].strip()[:70].ljust(70), st)
