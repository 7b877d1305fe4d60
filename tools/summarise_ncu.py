"""Turns the raw ncu outputs of a gpurun call into the small summaries committed under profiles/.
usage: python tools/summarise_ncu.py <round-tag> [gpurun_out]"""
import csv
import json
import os
import subprocess
import sys
from collections import defaultdict

tag = sys.argv[1]
src = sys.argv[2] if len(sys.argv) > 2 else "gpurun_out"
out = "profiles"
os.makedirs(out, exist_ok=True)

# ---- launch list: share of the step per kernel
rows = [r for r in csv.reader(open(os.path.join(src, f"launches_{tag}.csv"))) if len(r) > 5]
hdr = next(r for r in rows if "Kernel Name" in r)
ix = {h: i for i, h in enumerate(hdr)}
agg = defaultdict(lambda: [0, 0.0])
for r in rows:
    if r is hdr or not r[ix["ID"]].isdigit():
        continue
    name = r[ix["Kernel Name"]].split("(")[0]
    v = float(r[ix["Metric Value"]].replace(",", ""))
    unit = r[ix["Metric Unit"]]
    v *= {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(unit, 1e-6)
    agg[name][0] += 1
    agg[name][1] += v
tot = sum(v[1] for v in agg.values())
with open(os.path.join(out, f"{tag}_launches_summary.md"), "w") as f:
    f.write(f"# ncu launch list ({tag}): `bench.py --steps 2 --warmup 3 --no-e2e --no-cpu`, C3 100Mx256 k=1000, 1 B200\n\n")
    f.write("`ncu --metrics gpu__time_duration.sum --clock-control none` (cold-cache, serialised: compare SHARES)\n\n")
    f.write("| kernel | launches | total ms | share |\n|---|---:|---:|---:|\n")
    for name, (cnt, ms) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        f.write(f"| `{name}` | {cnt} | {ms:.3f} | {100 * ms / tot:.2f} % |\n")
    f.write(f"| total | {sum(v[0] for v in agg.values())} | {tot:.3f} | 100 % |\n")

# ---- full capture: key metrics per profiled launch
rep = os.path.join(src, f"prof_{tag}.ncu-rep")
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rr = list(csv.reader(raw.splitlines()))
h, units = rr[0], rr[1]
keys = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "launch__shared_mem_per_block_dynamic", "sm__cycles_elapsed.max", "smsp__inst_executed.sum"]
cols = [i for i, name in enumerate(h) if name in keys]
traffic = {}
traffic_ms = -1.0
with open(os.path.join(out, f"{tag}_ncu_full_summary.csv"), "w") as f:
    w = csv.writer(f)
    w.writerow([h[i] for i in cols])
    w.writerow([units[i] for i in cols])
    for r in rr[2:]:
        w.writerow([r[i] for i in cols])
        name = r[h.index("Kernel Name")]
        dur = float(r[h.index("gpu__time_duration.sum")])
        if ("assign_pair" in name or "assign_tc" in name) and dur > traffic_ms:   # the main (longest) assign launch
            traffic_ms = dur

            def val(metric):
                i = h.index(metric)
                mult = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1}.get(units[i], 1)
                return float(r[i]) * mult
            traffic = {"assign_dram_bytes_per_launch": val("dram__bytes_read.sum") + val("dram__bytes_write.sum"),
                       "rows": 100_000_000, "source": f"profiles/{tag}_ncu_full_summary.csv (ncu --set full)"}
if traffic:
    tp = os.path.join(out, "traffic.json")
    cur = json.load(open(tp)) if os.path.exists(tp) else {}
    cur["C3_n1"] = traffic
    json.dump(cur, open(tp, "w"), indent=1)
print(open(os.path.join(out, f"{tag}_launches_summary.md")).read())
print(open(os.path.join(out, f"{tag}_ncu_full_summary.csv")).read())
