"""Summarise an `ncu --page source --csv` export: stall totals and the hottest instructions."""
import csv
import sys

path, top = sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 25
rows = list(csv.reader(open(path)))
hdr = rows[1]
ia, isrc, isamp, iex = (hdr.index(x) for x in ("Address", "Source", "# Samples", "Instructions Executed"))
data, seen = [], set()
for r in rows[2:]:
    if len(r) > isamp and r[isamp].isdigit() and r[ia] not in seen:
        seen.add(r[ia])
        data.append(r)
stall = [(i, h) for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
tot = sum(int(r[isamp]) for r in data)
agg = {}
for r in data:
    for i, h in stall:
        if i < len(r) and r[i].isdigit():
            agg[h] = agg.get(h, 0) + int(r[i])
print("samples", tot, "instructions", len(data))
print(" ".join(f"{h[6:]}={v}" for h, v in sorted(agg.items(), key=lambda x: -x[1]) if v))
base = int(data[0][ia], 16)
for r in sorted(data, key=lambda r: -int(r[isamp]))[:top]:
    st = sorted([(int(r[i]), h[6:]) for i, h in stall if i < len(r) and r[i] not in ("", "0")], reverse=True)[:2]
    print((int(r[ia], 16) - base) // 16, r[isamp].rjust(5), r[iex].rjust(8), r[isrc].strip()[:66], st)
