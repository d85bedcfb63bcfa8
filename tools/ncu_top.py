"""Developer helper: top stalled SASS instructions from `ncu --page source --csv` output."""
import csv
import sys

path, top = sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 25
rows = list(csv.reader(open(path)))
start = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[start]
body = []
for r in rows[start + 1:]:
    if not r or r[0] in ("Kernel Name", "Address"):
        break
    body.append(r)
col = {h: i for i, h in enumerate(hdr)}
stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
tot = sum(int(r[col["# Samples"]] or 0) for r in body)
print("instructions:", len(body), "samples:", tot, "warp-inst executed:",
      sum(int(r[col["Instructions Executed"]] or 0) for r in body))
agg = {}
for r in body:
    for h in stall_cols:
        agg[h] = agg.get(h, 0) + int(r[col[h]] or 0)
print("stall totals:", {k: v for k, v in sorted(agg.items(), key=lambda kv: -kv[1]) if v})
body.sort(key=lambda r: -int(r[col["# Samples"]] or 0))
for r in body[:top]:
    n = int(r[col["# Samples"]] or 0)
    st = {h[6:]: int(r[col[h]] or 0) for h in stall_cols if int(r[col[h]] or 0)}
    st = dict(sorted(st.items(), key=lambda kv: -kv[1])[:3])
    print(f"{n:6d} {100 * n / max(tot, 1):5.1f}%  {r[col['Source']][:70]:70s} {st}")
