"""Warp-state samples of one captured kernel aggregated per SASS subroutine (the __noinline__ device functions are
emitted after the kernel body, each ending in RET.REL).  usage: python tools/ncu_by_subroutine.py file.ncu-rep
Capture with:  ncu --section SourceCounters --section WarpStateStats --sampling-interval 0 --import-source on ...  dev tool."""
import csv, subprocess, sys
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hi = [i for i, r in enumerate(rows) if "Address" in r][0]
h = rows[hi]
ia, isrc, isamp, iex = h.index("Address"), h.index("Source"), h.index("# Samples"), h.index("Instructions Executed")
data = [(r[isrc], int(r[isamp] or 0), int(r[iex] or 0)) for r in rows[hi + 1:] if len(r) > isamp and r[ia]]
tot = sum(d[1] for d in data)
print("total samples", tot, "instructions", len(data))
rets = [i for i, d in enumerate(data) if "RET.REL" in d[0]]
prev = 0
for r in rets + [len(data) - 1]:
    seg = data[prev:r + 1]
    if not seg:
        continue
    mix = {}
    for d in seg:
        t = d[0].split()
        op = (t[1] if t[0].startswith("@") else t[0]).split(".")[0]
        mix[op] = mix.get(op, 0) + 1
    top = sorted(mix.items(), key=lambda x: -x[1])[:4]
    s = sum(d[1] for d in seg)
    print("instr %5d..%5d (%5d)  samples %7d (%5.1f %%)  executed %9d  %s" % (prev, r, len(seg), s, 100.0 * s / max(tot, 1),
                                                                          sum(d[2] for d in seg), top))
    prev = r + 1
