"""Summarises an `ncu --set full` report into a markdown table per kernel and refreshes profiles/traffic.json.
usage: python tools/ncu_summary.py gpurun_out/prof.ncu-rep "<command that was profiled>" "config N" > profiles/<name>.md
(one section per distinct kernel name + template arguments + grid: the first captured instance)"""
import csv
import json
import os
import subprocess
import sys

WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sector_hit_rate.pct", "smsp__inst_executed.sum", "launch__grid_size", "launch__block_size",
        "launch__shared_mem_per_block_dynamic"]
UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}

rep, cmd = sys.argv[1], (sys.argv[2] if len(sys.argv) > 2 else "")
workload = sys.argv[3] if len(sys.argv) > 3 else "config 1"
seen = set()
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader([l for l in raw.splitlines() if not l.startswith("==")]))
hdr, units = rows[0], rows[1]
ki = hdr.index("Kernel Name")
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tpath = os.path.join(root, "profiles", "traffic.json")
traffic = json.load(open(tpath)) if os.path.exists(tpath) else {}
print("# `ncu --set full --clock-control none` capture\n")
if cmd:
    print("command: `%s` (after the same command exited 0 without ncu); report read with `ncu -i ... --page raw --csv`.\n" % cmd)
for r in rows[2:]:
    name = r[ki].split("(")[0].replace("void ", "").split("<")[0].split("::")[-1]
    ident = (r[ki].split("(")[0], r[hdr.index("launch__grid_size")] if "launch__grid_size" in hdr else "")
    if ident in seen:
        continue
    seen.add(ident)
    print("## %s\n\n`%s`\n\n| metric | value | unit |\n|---|---|---|" % (name, r[ki][:120]))
    stalls = {}
    for i, h in enumerate(hdr):
        if h in WANT:
            print("| %s | %s | %s |" % (h, r[i], units[i]))
        if h.startswith("smsp__pcsamp_warps_issue_stalled_") and not h.endswith("_not_issued"):
            try:
                stalls[h[len("smsp__pcsamp_warps_issue_stalled_"):]] = float(r[i].replace(",", ""))
            except ValueError:
                pass
    tot = sum(stalls.values()) or 1.0
    top = sorted(stalls.items(), key=lambda kv: -kv[1])[:6]
    print("| warp stall samples (top) | %s | share |" % ", ".join("%s %.0f%%" % (k, 100 * v / tot) for k, v in top))
    rd = float(r[hdr.index("dram__bytes_read.sum")].replace(",", "")) * UNIT[units[hdr.index("dram__bytes_read.sum")]]
    wr = float(r[hdr.index("dram__bytes_write.sum")].replace(",", "")) * UNIT[units[hdr.index("dram__bytes_write.sum")]]
    key = name if workload == "config 1" else "%s@%s" % (name, workload)
    if key not in traffic or not traffic[key].get("_fresh"):
        traffic[key] = {"dram_bytes_per_launch": rd + wr, "workload": workload, "source": "profiles/ (%s)" % os.path.basename(rep), "_fresh": True}
    print("\nDRAM traffic per launch: %.1f MB read + %.1f MB written = %.1f MB\n" % (rd / 1e6, wr / 1e6, (rd + wr) / 1e6))
traffic = {k: {kk: vv for kk, vv in v.items() if kk != "_fresh"} for k, v in traffic.items() if isinstance(v, dict) and v}
json.dump(traffic, open(tpath, "w"), indent=1)
