"""One table per `ncu --set full` report: duration, tensor-pipe activity, DRAM bytes and achieved GB/s against the
measured copy bandwidth, occupancy - the north_star's "evidence" view.
usage: python tools/ncu_evidence.py <report.ncu-rep> "<title>" [<report> "<title>" ...] > profiles/r2_ncu_evidence.md"""
import csv
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
TUNIT = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}
print("# ncu evidence, round 2 (`ncu --set full --clock-control none`, one capture per kernel instance, cold caches)\n")
print("HBM fraction = (dram__bytes_read.sum + dram__bytes_write.sum) / gpu__time_duration.sum against the measured copy "
      "bandwidth of %.0f GB/s (`MEASURED_PEAKS.json`); tensor pipe = sm__pipe_tensor_cycles_active.avg.pct_of_peak_"
      "sustained_active.  Kernels far below both ceilings are latency chains; DESIGN.md section 4 says which and why.\n" % peaks["hbm_gbs"])
args = sys.argv[1:]
for rep, title in zip(args[0::2], args[1::2]):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader([l for l in raw.splitlines() if not l.startswith("==")]))
    hdr, units = rows[0], rows[1]
    col = {h: i for i, h in enumerate(hdr)}
    print("## %s\n" % title)
    print("| kernel | grid x block | us | tensor pipe % | DRAM MB | GB/s | of HBM peak | warps active % | regs |")
    print("|---|---|---|---|---|---|---|---|---|")
    seen = set()
    for r in rows[2:]:
        name = r[col["Kernel Name"]].split("(")[0].replace("void ", "").replace("indnet::", "")
        key = (name, r[col["launch__grid_size"]])
        if key in seen:
            continue
        seen.add(key)
        f = lambda k: float(r[col[k]].replace(",", ""))  # noqa: E731
        us = f("gpu__time_duration.sum") * TUNIT[units[col["gpu__time_duration.sum"]]]
        mb = (f("dram__bytes_read.sum") * UNIT[units[col["dram__bytes_read.sum"]]] +
              f("dram__bytes_write.sum") * UNIT[units[col["dram__bytes_write.sum"]]]) / 1e6
        gbs = mb / 1e3 / (us * 1e-6)
        print("| `%s` | %s x %s | %.1f | %.1f | %.1f | %.0f | %.2f | %.0f | %s |" % (
            name[:60], r[col["launch__grid_size"]], r[col["launch__block_size"]], us,
            f("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"), mb, gbs, gbs / peaks["hbm_gbs"],
            f("sm__warps_active.avg.pct_of_peak_sustained_active"), r[col["launch__registers_per_thread"]]))
    print()
