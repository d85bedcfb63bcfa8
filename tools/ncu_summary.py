"""Developer helper: condense an .ncu-rep into the text summary committed under profiles/.

    python tools/ncu_summary.py gpurun_out/x.ncu-rep > profiles/x.summary.txt
"""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
WANT = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "launch__waves_per_multiprocessor",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sectors_srcunit_tex_op_read.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__inst_executed.sum",
    "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
    "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
col = {h: i for i, h in enumerate(hdr)}
print(f"# {rep}  (ncu --set full --clock-control none; kernels serialised, cold caches)")
for k, r in enumerate(rows[2:]):
    print(f"\n## launch {k}: {r[col['Kernel Name']][:150]}")
    for w in WANT:
        if w in col:
            print(f"{w:72s} {r[col[w]]:>16s} {units[col[w]]}")
    try:
        scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
        rd = float(r[col["dram__bytes_read.sum"]]) * scale[units[col["dram__bytes_read.sum"]]]
        wr = float(r[col["dram__bytes_write.sum"]]) * scale[units[col["dram__bytes_write.sum"]]]
        print(f"{'traffic = dram read + write':72s} {(rd + wr) / 1e6:16.3f} Mbyte")
    except Exception:
        pass
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
try:
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
    agg = {}
    for r in body:
        for h in stall_cols:
            agg[h] = agg.get(h, 0) + int(r[col[h]] or 0)
    print(f"\n## warp-stall samples, launch 0 (total {tot})")
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1]):
        if v:
            print(f"{k:28s} {v:7d} {100 * v / max(tot, 1):5.1f}%")
    print("\n## top instructions by samples")
    body.sort(key=lambda r: -int(r[col["# Samples"]] or 0))
    for r in body[:12]:
        n = int(r[col["# Samples"]] or 0)
        st = {h[6:]: int(r[col[h]] or 0) for h in stall_cols if int(r[col[h]] or 0)}
        st = dict(sorted(st.items(), key=lambda kv: -kv[1])[:2])
        print(f"{n:6d} {100 * n / max(tot, 1):5.1f}%  {r[col['Source']][:64]:64s} {st}")
except StopIteration:
    pass
