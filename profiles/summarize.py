"""Turn a Nsight Compute report into the short text summary kept under profiles/.

    python profiles/summarize.py gpurun_out/prof.ncu-rep > profiles/<name>.txt

Reads the report with `ncu -i ... --page raw --csv` (no GPU needed) and prints, per
profiled launch, the metrics the roofline discussion in DESIGN.md relies on.
"""
import csv
import io
import subprocess
import sys

WANT = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "launch__grid_size", "launch__block_size", "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic",
    "launch__waves_per_multiprocessor", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "smsp__cycles_active.avg", "sm__cycles_elapsed.max", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_alu.sum", "sm__inst_executed_pipe_fma.sum", "sm__inst_executed_pipe_fp64.sum",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "lts__t_sector_hit_rate.pct",
]


def main(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    name_i = hdr.index("Kernel Name")
    print(f"# {path}")
    for r in rows[2:]:
        print(f"\n## {r[name_i][:110]}")
        for w in WANT:
            if w in hdr:
                i = hdr.index(w)
                print(f"{w:70s} {r[i]} {units[i]}")
        if "dram__bytes_read.sum" in hdr:
            def num(metric):
                i = hdr.index(metric)
                v = float(r[i].replace(",", ""))
                u = units[i].lower()
                return v * {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}.get(u, 1)
            tr = num("dram__bytes_read.sum") + num("dram__bytes_write.sum")
            print(f"{'dram traffic per launch (read+write)':70s} {tr:.0f} byte")


if __name__ == "__main__":
    main(sys.argv[1])
