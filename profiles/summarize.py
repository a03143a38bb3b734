"""Turns the ncu outputs brought back in gpurun_out/ into the small text summaries committed here.
  python profiles/summarize.py launches gpurun_out/launches_X.csv > profiles/rN_X_launches.txt
  python profiles/summarize.py full gpurun_out/prof_X.ncu-rep   > profiles/rN_X_full.txt
"""
import collections
import csv
import io
import subprocess
import sys

METRICS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
           "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
           "sm__throughput.avg.pct_of_peak_sustained_elapsed",
           "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
           "sm__inst_executed_pipe_tensor.sum", "sm__warps_active.avg.pct_of_peak_sustained_active",
           "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
           "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_registers",
           "smsp__cycles_active.avg", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
           "lts__t_bytes.sum", "sm__cycles_elapsed.max"]


def launches(path):
    lines = [l for l in open(path) if not l.startswith("==")]
    agg = collections.defaultdict(lambda: [0, 0.0])
    for row in csv.DictReader(io.StringIO("".join(lines))):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(row["Metric Value"].replace(",", ""))
        v = {"ns": v / 1000, "us": v, "ms": v * 1000}.get(row["Metric Unit"], v)
        agg[row["Kernel Name"][:110]][0] += 1
        agg[row["Kernel Name"][:110]][1] += v
    tot = sum(v[1] for v in agg.values())
    print(f"# {path}: per-kernel device time over the captured launches (cold-cache, serialised: compare SHARES)")
    print(f"# total {tot:.1f} us over {sum(v[0] for v in agg.values())} launches")
    for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{t:10.1f} us {100 * t / tot:5.1f}%  n={n:4d}  avg {t / n:8.2f} us  {k}")


def full(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    print(f"# {path}: ncu --set full, selected metrics per captured launch")
    for row in rows[2:]:
        print("kernel:", row[hdr.index("Kernel Name")][:120])
        for m in METRICS:
            if m in hdr:
                i = hdr.index(m)
                print(f"    {m:70s} {row[i]:>16s} {units[i]}")


if __name__ == "__main__":
    {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2])
