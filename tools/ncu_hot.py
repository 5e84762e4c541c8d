"""Top stall sites of one kernel in an ncu report (SASS view with stall samples).
usage: python tools/ncu_hot.py report.ncu-rep <kernel-name-regex> [N]"""
import csv
import subprocess
import sys

rep, kern = sys.argv[1], sys.argv[2]
topn = int(sys.argv[3]) if len(sys.argv) > 3 else 25
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + kern],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hi = next(i for i, r in enumerate(rows) if len(r) > 2 and r[0] == "Address")
hdr = rows[hi]
si, src = hdr.index("# Samples"), hdr.index("Source")
stall = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
data = []
for r in rows[hi + 1:]:
    if len(r) <= si or r[0] == "Address" or r[0] == "Kernel Name":
        if r and r[0] == "Kernel Name":
            break
        continue
    try:
        v = float(r[si].replace(",", ""))
    except ValueError:
        continue
    why = sorted(((float(r[i] or 0), hdr[i][6:]) for i in stall), reverse=True)[:2]
    data.append((v, r[src][:100], why))
tot = sum(d[0] for d in data) or 1.0
for k, (v, s, why) in enumerate(data):
    data[k] = (v, k, s, why)
for v, k, s, why in sorted(data, reverse=True)[:topn]:
    print("%5.1f%%  #%-5d %-100s %s" % (100 * v / tot, k, s, ", ".join("%s %.0f" % (n, x) for x, n in why if x)))
