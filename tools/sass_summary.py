"""Per kernel of gnn-cca_b200/libgnncca.so: instruction count and the counts of the SASS mnemonics that identify the Blackwell paths.

    python tools/sass_summary.py > profiles/r02_sass_summary.txt      (runs on the CPU box: cuobjdump only)
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "gnn-cca_b200", "libgnncca.so")
KEEP = re.compile(r"^(UTC\w*MMA|UTCBAR|LDTM|STTM|UTMALDG|UTMASTG|UBLKCP|HMMA|IMMA|FFMA2|FADD2|DFMA|SHFL|RED|REDG|REDUX|ATOMG|ATOMS|ATOM)\b")


def main():
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    print("# SASS evidence per kernel of gnn-cca_b200/libgnncca.so (cuobjdump -sass, sm_100a): counts of the tensor / TMA / bulk-copy / packed-fp32 / atomic mnemonics")
    print("# tcgen05.mma -> UTC*MMA, tcgen05.ld/st -> LDTM/STTM, cp.async.bulk.tensor -> UTMALDG, cp.async.bulk -> UBLKCP, mma.sync -> HMMA, fma.rn.f32x2 -> FFMA2; RED/ATOMG = global atomics")
    print("# regenerate: python tools/sass_summary.py\n")
    name, counts, n = None, collections.Counter(), 0

    def flush():
        if name is not None:
            tags = " ".join("%s:%d" % kv for kv in sorted(counts.items()))
            print("%-110s instrs=%5d  %s" % (name[:110], n, tags))
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            flush()
            name, counts, n = re.sub(r"^_ZN?6?gnncca", "", m.group(1)), collections.Counter(), 0
            continue
        m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)", line)
        if m:
            n += 1
            op = m.group(1)
            if KEEP.match(op):
                counts[op.split(".")[0] if op.startswith(("SHFL", "FFMA2", "FADD2", "DFMA", "HMMA")) else op] += 1
    flush()


if __name__ == "__main__":
    sys.exit(main())
