"""Developer helper: per-kernel totals from an `ncu --metrics gpu__time_duration.sum --csv` launch list."""
import csv, sys
from collections import defaultdict
path, title = sys.argv[1], sys.argv[2]
rows = [r for r in csv.reader(open(path)) if len(r) > 10]
hdr = rows[0]
ci = {h: i for i, h in enumerate(hdr)}
agg = defaultdict(lambda: [0, 0.0])
for r in rows[1:]:
    name = r[ci["Kernel Name"]]
    name = name.split("(")[0].replace("void ", "").replace("<unnamed>::", "")[:70]
    agg[name][0] += 1
    agg[name][1] += float(r[ci["Metric Value"]])
tot = sum(v[1] for v in agg.values())
print(f"# {title}")
print("# ncu --metrics gpu__time_duration.sum --clock-control none (per-launch times are cold-cache and serialised: read the SHARES)")
print("# kernel, launches, total us, avg us, share")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{k}, {v[0]}, {v[1]/1e3:.1f}, {v[1]/1e3/v[0]:.2f}, {100*v[1]/tot:.1f}%")
