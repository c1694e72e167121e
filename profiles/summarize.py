"""Turns the ncu exports under gpurun_out/ into the small tracked summaries in profiles/.
    python profiles/summarize.py launches gpurun_out/r1_launches_ecoli.csv > profiles/r1_launches_ecoli_summary.txt
    python profiles/summarize.py raw gpurun_out/r1_prof_ecoli.ncu-rep > profiles/r1_ncu_full_ecoli.txt
    python profiles/summarize.py traffic gpurun_out/r2_launches_human.csv human     # -> profiles/ncu_traffic.json
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


# kernel name -> bench.py kernel class (the ProfScope classes of csrc/gtsfx_b200.cu)
CLASS_OF = [("k_count_dna", "count"), ("k_tile_subsort", "sort_lcp"), ("k_kmers<2, 0>", "count"), ("k_kmers<2, 2>", "count"), ("k_kmers<8, 0>", "count"), ("k_kmers<8, 2>", "count"),
            ("k_kmers<2, 1>", "scatter"), ("k_kmers<8, 1>", "scatter"), ("k_keyed_partition", "scatter"),
            ("k_stage_partition", "scatter"), ("k_partition", "scatter"), ("k_scatter_", "scatter"), ("k_bins_init", "scatter"),
            ("k_keyed_chunk_table", "scatter"), ("k_chunk_table", "scatter"), ("k_tile_sort", "sort_lcp"),
            ("k_sort_buckets", "sort_lcp"), ("k_sort_medium", "sort_medium"), ("k_boundary", "boundary"),
            ("k_part_first", "boundary"), ("k_widen", "widen"), ("k_tail", "tail"), ("k_part_init", "part_init"),
            ("k_padded", "padded"), ("k_place_padded", "padded")]


def traffic(path, workload, out="profiles/ncu_traffic.json"):
    """dram bytes (read + write) per step and kernel class from a launch list taken with
    --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum; steps = launches of k_part_init."""
    import json
    import os
    lines = [l for l in open(path) if not l.startswith("==")]
    per = collections.defaultdict(float)
    steps = 0
    for row in csv.DictReader(lines):
        name = row["Kernel Name"]
        if row["Metric Name"] == "gpu__time_duration.sum" and name.startswith("k_part_init"):
            steps += 1
        if not row["Metric Name"].startswith("dram__bytes"):
            continue
        unit = row.get("Metric Unit", "byte")
        scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1.0)
        for key, cls in CLASS_OF:
            if key in name:
                per[cls] += float(row["Metric Value"].replace(",", "")) * scale
                break
    db = json.load(open(out)) if os.path.exists(out) else {}
    db[workload] = {k: v / max(steps, 1) for k, v in per.items()}
    db[workload]["source"] = os.path.basename(path)
    db[workload]["steps_in_capture"] = steps
    json.dump(db, open(out, "w"), indent=1, sort_keys=True)
    print(json.dumps(db[workload], indent=1))


def source(path, kernel_regex, top=30):
    """Instructions executed and stall samples per CUDA source line of one kernel (needs -lineinfo, --import-source on)."""
    out = subprocess.run(["ncu", "-i", path, "--page", "source", "--csv", "--print-source", "cuda,sass", "-k",
                          "regex:" + kernel_regex], capture_output=True, text=True).stdout
    cur, hdr = None, None
    agg, samp, src = collections.Counter(), collections.Counter(), {}
    for r in csv.reader(out.splitlines()):
        if not r:
            continue
        if r[0] == "File Path":
            cur = r[1].split("/")[-1]
            continue
        if r[0] == "Line No":
            hdr = r
            continue
        if r[0] == "Function Name" or hdr is None:
            continue
        try:
            ln = int(r[0])
        except ValueError:
            continue
        if r[2] in ("-", ""):
            agg[(cur, ln)] += int(r[hdr.index("Instructions Executed")] or 0)
            samp[(cur, ln)] += int(r[hdr.index("# Samples")] or 0)
            src[(cur, ln)] = r[1].strip()[:100]
    tot, ts = sum(agg.values()), sum(samp.values())
    print(f"== {kernel_regex}: {tot} warp instructions, {ts} stall samples")
    for k, v in agg.most_common(int(top)):
        print(f"{v / max(tot, 1) * 100:5.1f}% inst {samp[k] / max(ts, 1) * 100:5.1f}% samples  {k[0]}:{k[1]}  {src[k]}")


if __name__ == "__main__":
    if sys.argv[1] == "source":
        source(sys.argv[2], sys.argv[3], *(sys.argv[4:5]))
    elif sys.argv[1] == "traffic":
        traffic(sys.argv[2], sys.argv[3])
    else:
        {"launches": launches, "raw": raw}[sys.argv[1]](sys.argv[2])
