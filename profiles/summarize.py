"""Turns the ncu exports under gpurun_out/ into the small tracked summaries in profiles/.
    python profiles/summarize.py launches gpurun_out/r1_launches_ecoli.csv > profiles/r1_launches_ecoli_summary.txt
    python profiles/summarize.py raw gpurun_out/r1_prof_ecoli.ncu-rep > profiles/r1_ncu_full_ecoli.txt
"""
import collections
import csv
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_bytes.sum",
        "smsp__inst_executed.sum", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum"]


def launches(path):
    lines = [l for l in open(path) if not l.startswith("==")]
    agg = collections.OrderedDict()
    for row in csv.DictReader(lines):
        k = row["Kernel Name"].split("(")[0][:44]
        a = agg.setdefault(k, collections.defaultdict(float))
        a[row["Metric Name"]] += float(row["Metric Value"].replace(",", ""))
        a["n_" + row["Metric Name"]] += 1
    tot = sum(a["gpu__time_duration.sum"] for a in agg.values())
    print(f"{'kernel':44s} {'n':>4s} {'us/launch':>10s} {'share':>6s} {'dram rd MB':>11s} {'dram wr MB':>11s}")
    for k, a in sorted(agg.items(), key=lambda x: -x[1]["gpu__time_duration.sum"]):
        n = a["n_gpu__time_duration.sum"]
        print(f"{k:44s} {int(n):4d} {a['gpu__time_duration.sum']/n/1e3:10.1f} {a['gpu__time_duration.sum']/tot:6.3f} "
              f"{a['dram__bytes_read.sum']/n/1e6:11.1f} {a['dram__bytes_write.sum']/n/1e6:11.1f}")


def raw(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        print("==", r[hdr.index("Kernel Name")][:100])
        for k in KEYS:
            if k in hdr:
                print(f"   {k} = {r[hdr.index(k)]} {units[hdr.index(k)]}")


if __name__ == "__main__":
    {"launches": launches, "raw": raw}[sys.argv[1]](sys.argv[2])
