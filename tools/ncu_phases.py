"""Cumulative stall samples between the synchronisation points of one kernel (SASS order) - shows which phase of a
barrier-separated kernel the warps spend their time in.  usage: python tools/ncu_phases.py report.ncu-rep <kernel-regex>"""
import csv, subprocess, sys
rep, kern = sys.argv[1], sys.argv[2]
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + kern], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hi = next(i for i, r in enumerate(rows) if len(r) > 2 and "# Samples" in r)
hdr = rows[hi]; si = hdr.index("# Samples"); src = hdr.index("Source")
data = []
for r in rows[hi + 1:]:
    try: data.append((float(r[si].replace(",", "")), r[src]))
    except (ValueError, IndexError): pass
tot = sum(d[0] for d in data) or 1
acc = last = 0
KEYS = ("BAR.SYNC", "SYNCS.PHASECHK", "UTCHMMA", "LDTM", "UTMALDG", "UTMASTG", "LDGDEPBAR", "DEPBAR", "EXIT")
prev = None
for i, (v, s) in enumerate(data):
    acc += v
    k = next((k for k in KEYS if k in s), None)
    if k and (k != prev or acc - last > 0.01 * tot):
        print("%5d  cum %5.1f%%  (+%4.1f%%)  %s" % (i, 100 * acc / tot, 100 * (acc - last) / tot, s[:80])); last = acc
    if k: prev = k
