"""Per-kernel counts of the Blackwell-specific SASS mnemonics in libindnet.so (proof that the tcgen05 / TMEM / TMA
path is what was compiled).  usage: python tools/sass_summary.py > profiles/r2_sass_summary.md"""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "induction-network-on-fewrel_b200", "csrc", "libindnet.so")
WANT = ["UTCHMMA", "UTMALDG", "UTMASTG", "UBLKCP", "LDTM", "STTM", "UTCBAR", "UTCATOMSWS", "SYNCS", "MUFU.TANH", "UCGABAR", "ATOMS", "LDGSTS"]
out = subprocess.run(["cuobjdump", "-sass", SO], capture_output=True, text=True).stdout
counts = collections.OrderedDict()
cur = None
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0]
        cur = cur.replace("void ", "").replace("indnet::", "")
        counts.setdefault(cur, collections.Counter())
        continue
    if cur:
        for w in WANT:
            if re.search(r"\b" + re.escape(w), line):
                counts[cur][w] += 1
print("# SASS summary of `csrc/libindnet.so` (sm_100a)\n")
print("command: `cuobjdump -sass induction-network-on-fewrel_b200/csrc/libindnet.so` (nvcc 12.9, `-gencode arch=compute_100a,code=sm_100a`), "
      "counted per kernel by `tools/sass_summary.py`.\n")
print("`UTCHMMA` = tcgen05.mma, `UTMALDG` / `UTMASTG` = TMA tile load / store, `UBLKCP` = cp.async.bulk (incl. distributed "
      "shared memory), `LDTM` / `STTM` = tcgen05.ld / st (tensor memory), `UTCBAR` = tcgen05.commit, `UTCATOMSWS` = TMEM "
      "alloc / dealloc, `SYNCS` = mbarrier, `UCGABAR` = cluster barrier.\n")
print("| kernel | " + " | ".join(WANT) + " |\n|---|" + "---|" * len(WANT))
for k, c in counts.items():
    if any(c[w] for w in WANT[:8]):
        print("| `%s` | %s |" % (k[:70], " | ".join(str(c[w]) if c[w] else "" for w in WANT)))
tot = collections.Counter()
for c in counts.values():
    tot.update(c)
print("| **all kernels (%d)** | %s |" % (len(counts), " | ".join(str(tot[w]) for w in WANT)))
