"""Turns gpurun_out ncu artefacts into the small text summaries committed under profiles/.

usage: python profiles/summarize.py list  gpurun_out/launches_X.csv         > profiles/X_launches.txt
       python profiles/summarize.py full  gpurun_out/prof_X.ncu-rep [regex] > profiles/X_full.txt
"""
import collections
import csv
import io
import subprocess
import sys

METRICS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sectors_srcunit_tex_op_read.sum",
    "lts__t_sector_hit_rate.pct", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__t_sector_hit_rate.pct",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_tensor.sum", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__shared_mem_per_block_dynamic",
    "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
]


def launches(path):
    lines = [l for l in open(path) if not l.startswith("==")]
    agg = collections.defaultdict(lambda: [0, 0.0])
    n = 0
    for row in csv.DictReader(io.StringIO("".join(lines))):
        v = float(row["Metric Value"].replace(",", ""))
        v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(row["Metric Unit"], 1.0)
        agg[row["Kernel Name"]][0] += 1
        agg[row["Kernel Name"]][1] += v
        n += 1
    tot = sum(v[1] for v in agg.values())
    print(f"# {path}: {n} launches, {tot / 1e3:.2f} ms of kernel time (ncu: cold cache, serialised; compare SHARES)")
    print(f"{'share':>7} {'total ms':>10} {'launches':>8}  kernel")
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{v[1] / tot * 100:6.2f}% {v[1] / 1e3:10.3f} {v[0]:8d}  {k[:110]}")


def full(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    print(f"# {path}: ncu --set full --clock-control none, per captured launch")
    for r in rows[2:]:
        print(f"--- launch id {r[idx['ID']]}: {r[idx['Kernel Name']][:100]}")
        for m in METRICS:
            if m in idx:
                print(f"    {m:68s} {r[idx[m]]:>20s} {units[idx[m]]}")


if __name__ == "__main__":
    {"list": launches, "full": full}[sys.argv[1]](sys.argv[2])
