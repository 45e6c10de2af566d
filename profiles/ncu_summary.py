"""Text summary of one .ncu-rep (first kernel in it): headline metrics + per-source-line warp-stall samples.
python profiles/ncu_summary.py X.ncu-rep > profiles/NAME.txt"""
import csv
import io
import subprocess
import sys
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
h, v = rows[0], rows[-1]
print("#", rep)
print("##", v[h.index("Kernel Name")] if "Kernel Name" in h else "")
want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__cycles_elapsed.max", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "lts__t_sector_hit_rate.pct"]
units = rows[1] if len(rows) > 2 else [""] * len(h)
for w in want:
    if w in h:
        i = h.index(w)
        print(f"{w:70s} {v[i]} {units[i]}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"], capture_output=True, text=True).stdout
open("/tmp/_src.csv", "w").write(src)
print("\n## warp-stall samples per source line (>= 1.5 % of all samples)")
print(subprocess.run([sys.executable, __file__.replace("ncu_summary.py", "summarize_source.py"), "/tmp/_src.csv", "0.015"], capture_output=True, text=True).stdout)
