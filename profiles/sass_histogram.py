"""SASS opcode histogram per kernel of the built library (the mnemonics that prove tcgen05 / TMA / packed math):
python profiles/sass_histogram.py > profiles/r2_sass_histogram.txt"""
import collections
import os
import re
import subprocess
so = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "k-means-demo_b200", "libkmeans.so")
out = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
WATCH = ["UTCHMMA", "UTCBAR", "LDTM", "UTMALDG", "UTMALDG.2D.GATHER4", "SYNCS", "FADD2", "FFMA2", "FMUL2", "FMNMX3", "REDG", "ATOMS", "ATOMG", "RED",
         "LDG", "STG", "LDS", "STS", "DADD", "DMUL", "DFMA", "F2FP", "SHFL", "VOTE", "REDUX", "ELECT", "NANOSLEEP", "MEMBAR", "BAR"]
cur, hist = None, collections.OrderedDict()
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0]
        hist[cur] = collections.Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m and cur:
        op = m.group(1)
        hist[cur]["total"] += 1
        if op.startswith("UTMALDG") and "GATHER4" in op:
            hist[cur]["UTMALDG.2D.GATHER4"] += 1
        for w in WATCH:
            if w != "UTMALDG.2D.GATHER4" and op.split(".")[0] == w:
                hist[cur][w] += 1
for k, c in hist.items():
    if c["total"] < 50:
        continue
    print(f"{k}: {c['total']} instructions")
    print("   " + "  ".join(f"{w}={c[w]}" for w in WATCH if c[w]))
