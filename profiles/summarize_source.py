"""Per-source-line warp-stall samples from `ncu -i X.ncu-rep --page source --print-source cuda,sass --csv`:
python profiles/summarize_source.py file.csv [min_share]"""
import csv
import sys
rows = list(csv.reader(open(sys.argv[1])))
thr = float(sys.argv[2]) if len(sys.argv) > 2 else 0.01
cur, hdr, out, tot = None, None, [], 0
for r in rows:
    if len(r) == 2 and r[0] == "File Path":
        cur = r[1].split("/")[-1]
    elif len(r) > 6 and r[0] == "Line No":
        hdr = {n: i for i, n in enumerate(r)}
        si = [i for i, n in enumerate(r) if n == "# Samples"][0]
        ei = [i for i, n in enumerate(r) if n == "Instructions Executed"][0]
    elif hdr and len(r) > 6 and r[0].isdigit() and r[2] == "-":   # a source line row (SASS rows carry an address)
        s = int(r[si]); tot += s
        out.append((s, int(r[ei]), cur, int(r[0]), r[1].strip()[:100]))
print("total samples", tot)
for s, e, f, ln, src in sorted(out, reverse=True):
    if s >= thr * tot:
        print(f"{s:7d} {100.0 * s / tot:5.1f}%  inst {e:10d}  {f}:{ln}  {src}")
