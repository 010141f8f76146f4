"""Summarise an ncu report's SASS page for one kernel: instruction / stall-sample share of every code segment
between BAR.SYNCs and the opcode mix.  Usage: python tools/ncu_segments.py report.ncu-rep kernel_regex"""
import collections
import csv
import subprocess
import sys

rep, pat = sys.argv[1], sys.argv[2]
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "-k", "regex:" + pat, "--launch-count", "1"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = next(r for r in rows if r and r[0] == "Address")
ci = {n: i for i, n in enumerate(hdr)}
S, I, X = ci["# Samples"], ci["Instructions Executed"], ci["L1 Wavefronts Shared Excessive"]


def iv(x):
    try:
        return int(x)
    except ValueError:
        return 0


data = [r for r in rows if r and r[0].startswith("0x") and len(r) > X]
ts, ti = sum(iv(r[S]) for r in data), sum(iv(r[I]) for r in data)
print("warp instructions", ti, "samples", ts)
seg, segs = 0, collections.OrderedDict()
for r in data:
    if "BAR.SYNC" in r[1]:
        seg += 1
    d = segs.setdefault(seg, [0, 0, 0, collections.Counter()])
    d[0] += iv(r[S]); d[1] += iv(r[I]); d[2] += iv(r[X])
    d[3][r[1].split()[0].split(".")[0] if not r[1].strip().startswith("@") else r[1].split()[1].split(".")[0]] += iv(r[I])
for k, v in segs.items():
    if v[1] * 200 < ti:
        continue
    print("segment %2d: instr %5.1f%%  samples %5.1f%%  excess smem wavefronts %d  top ops: %s" %
          (k, 100 * v[1] / ti, 100 * v[0] / ts, v[2], ", ".join("%s %.0f%%" % (o, 100 * c / v[1]) for o, c in v[3].most_common(7))))
