#!/usr/bin/env python
"""Benchmark of the enhanced-suffix-array path (`gt suffixerator -suf -lcp -bck`) on B200.

    python bench.py --gpus N --steps K --warmup W [--workload ecoli|chr1|human|protein] [--impl reference]

A step = one pass of the whole hot path over one synthetic genome: k-mer counting + bucket prefix
sums, insertion, in-bucket sort, LCP, tail.  metric = suffixes/s (suffixes = totallength + 1).

 * value : device time of K steps with the packed sequence already resident in HBM (CUDA events on
           the library's stream, L2 flushed between steps, max over ranks).
 * e2e   : the same work through the host-facing call (Sfxiterator over the C ABI) with HOST buffers:
           pinned H2D of the 2-bit sequence + range table, D2H of .suf (8 B/suffix) + .lcp + .bck tables.
 * N > 1 : one process per GPU.  The genome grows with N (N x the single-GPU size: weak scaling; --scaling strong
           keeps the named genome and splits it, which is how the 3.1 Gbp target runs on 8 GPUs);
           rank 0 uploads, the packed sequence is broadcast over NCCL, every rank counts (replicated,
           n/4 bytes) and sorts + LCP-annotates its own bucket range (the reference's -parts split).
           No collective inside the timed device region.
 * --impl reference : the UNMODIFIED reference (oracle/_ref/gt_ref, single thread because -lcp forbids
           -j > 1, src/match/sfx-opt.c:120-126) on the same workload, rank 0 only.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from genometools_b200 import encseq, synth  # noqa: E402

WORKLOADS = {
    # name: (per-GPU total length, description)
    "ecoli": (4_600_000, "configs[1]: synthetic 4.6 Mbp single-sequence random DNA (seed 42)"),
    "chr1": (250_000_000, "configs[2]: synthetic 250 Mbp DNA, 24 sequences, N-runs (seed 250)"),
    "human": (3_100_000_000, "configs[3]: synthetic 3.1 Gbp multi-FASTA DNA (seed 31)"),
    "protein": (1_000_000_000, "configs[4]: synthetic 1 G-residue protein set (seed 5)"),
    "genomelike": (50_000_000, "stress (not in BASELINE.json): 50 Mbp with repeat families, segmental duplications, N-runs"),
}


SHRINK = 1   # --shrink D: exploratory runs on 1/D of the named size (the line says so in config)


def make_workload(name: str, scale: int):
    per_gpu, _ = WORKLOADS[name]
    total = per_gpu * scale // SHRINK
    if name == "ecoli":
        return synth.random_dna(total, seed=42), "dna"
    if name == "chr1":
        return synth.multi_dna_with_nruns(total, nseq=24, seed=250), "dna"
    if name == "human":
        return synth.multi_dna_with_nruns(total, nseq=24, seed=31), "dna"
    if name == "genomelike":
        return synth.genome_like_dna(total, seed=77), "dna"
    return synth.random_protein(max(1, total // 334), seed=5), "protein"


def ncu_traffic(workload: str, kclass: str):
    """dram__bytes_read.sum + dram__bytes_write.sum per step of one kernel class, as profiles/summarize.py extracted it
    from the ncu launch list of this same command (profiles/ncu_traffic.json: {workload: {class: bytes}}).
    None when no capture of the current kernels exists - never a typed-in constant."""
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as fh:
            v = json.load(fh).get(workload, {}).get(kclass)
        return float(v) if v is not None else None
    except Exception:
        return None


def full_size_reference(workload: str):
    """Whole-workload time of the unmodified reference (one thread), recorded when tests/golden/scale_fingerprints.json
    was generated in the build container - context for the bounded sample timed live on this box."""
    try:
        with open(os.path.join(ROOT, "tests", "golden", "scale_fingerprints.json")) as fh:
            rec = json.load(fh)[workload]
        t = rec["reference"]["times_s"]
        path_s = sum(t.get(k, 0.0) for k in ("counting prefix distribution", "inserting suffixes into buckets",
                                              "sorting the buckets"))
        return {"where": "build container (not this box), 1 thread", "overall_s": t.get("overall"),
                "count_insert_sort_s": round(path_s, 2), "suffixes": rec["totallength"] + 1}
    except Exception:
        return None


def peak_hbm():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as fh:
            return float(json.load(fh)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, device: int):
        self.path = tempfile.mktemp(prefix="clocks_", suffix=".csv")
        self.proc = None
        try:
            self.fh = open(self.path, "w")
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(device), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=self.fh, stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        self.fh.close()
        sm, mx, reasons = [], [], set()
        for line in open(self.path):
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        os.unlink(self.path)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def write_fasta_for(enc, alphabet, path):
    encseq.write_fasta(path, encseq.decode_to_fasta_bytes(enc, alphabet))


def run_reference_once(fasta, alphabet):
    from oracle import oracle
    ref = oracle.run_reference(fasta, alphabet, showtime=True)
    t = ref["times"]
    path_s = sum(t.get(k, 0.0) for k in ("counting prefix distribution", "inserting suffixes into buckets",
                                          "sorting the buckets"))
    return ref["wall_s"], t.get("overall", ref["wall_s"]), path_s


def time_dropin(fasta, alphabet, tmp):
    gt_b200 = os.path.join(ROOT, "oracle", "_ref", "gt_b200")
    if not os.path.exists(gt_b200):
        return None
    cmd = [gt_b200, "suffixerator", "-" + alphabet, "-suf", "-lcp", "-bck", "-tis", "-indexname", os.path.join(tmp, "b200"),
           "-db", fasta]
    best = None
    for _ in range(2):                      # first run pays CUDA context creation and page-locking
        t0 = time.perf_counter()
        p = subprocess.run(cmd, capture_output=True, text=True)
        dt = time.perf_counter() - t0
        if p.returncode != 0:
            return {"error": p.stderr[-200:]}
        best = dt if best is None else min(best, dt)
    return {"gt_b200_wall_s": round(best, 3), "includes": "process start, CUDA context, FASTA parse, encoding, index files"}


def cpu_baseline(enc, alphabet, budget_n=12_000_000):
    """The reference's own CPU implementation on a bounded sample of the workload (rank 0, N=1)."""
    from oracle import oracle
    n = enc.shape[0]
    sample = enc if n <= budget_n else enc[:budget_n]
    what = "whole workload" if n <= budget_n else f"first {budget_n} symbols of the workload"
    if oracle.have_reference():
        with tempfile.TemporaryDirectory() as tmp:
            fa = os.path.join(tmp, "sample.fna")
            write_fasta_for(sample, alphabet, fa)
            wall, overall, path_s = run_reference_once(fa, alphabet)
            dropin = time_dropin(fa, alphabet, tmp)
        secs = max(overall, 1e-9)
        return {"value": (sample.shape[0] + 1) / secs, "unit": "suffixes/s", "cores": 1, "kind": "reference",
                "sample": f"{what}; oracle/_ref/gt_ref suffixerator -suf -lcp -bck -tis, 1 thread (-lcp forbids -j>1), "
                          f"TIME overall {overall:.2f} s of which count+insert+sort {path_s:.2f} s, wall {wall:.2f} s",
                # the hot path alone (what `value` covers): count + insert + sort phases of the same run
                "path_only_value": (sample.shape[0] + 1) / max(path_s, 1e-9),
                # where a `gt` user sits: the reference binary with only the Sfxiterator swapped for the C ABI
                # (oracle/_ref/gt_b200), same FASTA, same options, wall clock incl. FASTA parsing, encoding, file output
                "dropin_same_sample": dropin,
                "host_cores_available": os.cpu_count()}
    small = sample[:2_000_000]
    t0 = time.perf_counter()
    oracle.build(small, encseq.alphabet_size(alphabet))
    secs = time.perf_counter() - t0
    return {"value": (small.shape[0] + 1) / secs, "unit": "suffixes/s", "cores": 1, "kind": "port",
            "sample": f"first {small.shape[0]} symbols, oracle/sfx_oracle.c (reference binary not built)",
            "host_cores_available": os.cpu_count()}


def bench_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import oracle
    enc, alphabet = make_workload(args.workload, 1 if args.scaling == "strong" else args.gpus)
    budget = 12_000_000
    sample = enc if enc.shape[0] <= budget else enc[:budget]
    what = "whole workload" if sample is enc else f"first {budget} symbols"
    times = []
    kind = "reference" if oracle.have_reference() else "port"
    with tempfile.TemporaryDirectory() as tmp:
        fa = os.path.join(tmp, "w.fna")
        write_fasta_for(sample, alphabet, fa)
        for i in range(args.warmup + args.steps):
            if kind == "reference":
                _, overall, _ = run_reference_once(fa, alphabet)
            else:
                t0 = time.perf_counter()
                oracle.build(sample, encseq.alphabet_size(alphabet))
                overall = time.perf_counter() - t0
            if i >= args.warmup:
                times.append(overall)
    per = sum(times) / len(times)
    value = (sample.shape[0] + 1) / per
    line = {
        "impl": "reference", "metric": "suffixes/sec for SA+LCP", "value": value, "unit": "suffixes/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": per * 1e3,
        "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "u64", "data": "synthetic",
        "config": {"workload": WORKLOADS[args.workload][1], "totallength": int(enc.shape[0]), "sample": what},
        "cpu_baseline": {"value": value, "unit": "suffixes/s", "cores": 1, "kind": kind,
                         "sample": f"{what}; gt suffixerator -suf -lcp -bck -tis, single thread (the reference "
                                   f"refuses -lcp with -j>1), TIME overall per run"},
        "e2e": {"value": value, "unit": "suffixes/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    # BASELINE.json quotes the metric on the 3.1 Gbp genome (configs[3]); it fits one B200
    ap.add_argument("--workload", default="human", choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--e2e-steps", type=int, default=None)
    ap.add_argument("--skip-e2e", action="store_true", help="exploratory runs only: leaves e2e out of the line")
    ap.add_argument("--scaling", default="strong", choices=["weak", "strong"],
                    help="strong (default): the named genome is split over the N GPUs by bucket range (BASELINE.json's "
                         "3.1 Gbp at 1/2/4/8 GPUs); weak: genome = N x the named size (must stay below 2^32 symbols)")
    ap.add_argument("--verify", action="store_true", help="N > 1: compare every rank's part with a single-context run on rank 0")
    ap.add_argument("--shrink", type=int, default=1, help="exploratory: run on 1/D of the named workload size")
    args = ap.parse_args()
    global SHRINK
    SHRINK = max(1, args.shrink)
    if os.environ.get("NCCL_DEBUG", "").upper() == "VERSION":
        os.environ["NCCL_DEBUG"] = "WARN"        # keep stdout to the one JSON line
    if args.impl == "reference":
        bench_reference(args)
        return

    import ctypes as C
    import torch
    import torch.distributed as dist
    from genometools_b200 import sfx

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torch.distributed.run")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    if world > 1:
        # one PCIe link per rank: the host cores that would widen 32-bit positions are the narrower resource then
        # (the library's own multi-GPU handle makes the same choice, csrc/multi.cuh)
        os.environ.setdefault("GTB200_NARROW", "0")
    L = sfx.load_library()
    alphabet = "protein" if args.workload == "protein" else "dna"
    sigma = encseq.alphabet_size(alphabet)

    # ---- inputs: rank 0 synthesises, packs and pins; NCCL broadcast puts them in every GPU's HBM ----
    meta = torch.zeros(3, dtype=torch.int64, device=dev)
    enc = None
    if rank == 0:
        enc, _ = make_workload(args.workload, 1 if args.scaling == "strong" else world)
        n = int(enc.shape[0])
        if alphabet == "dna":
            words_h = torch.from_numpy(encseq.twobit_pack(enc).view(np.int64)).pin_memory()
            ranges_h = torch.from_numpy(encseq.special_ranges(enc).astype(np.uint64).view(np.int64).reshape(-1)).pin_memory()
        else:
            words_h = torch.from_numpy(enc).pin_memory()
            ranges_h = torch.zeros(0, dtype=torch.int64).pin_memory()
        meta = torch.tensor([n, words_h.numel(), ranges_h.numel()], dtype=torch.int64, device=dev)
    if world > 1:
        dist.broadcast(meta, 0)
    n, nwords, nrangewords = (int(x) for x in meta.tolist())
    wdtype = torch.int64 if alphabet == "dna" else torch.uint8
    words_d = torch.empty(nwords, dtype=wdtype, device=dev)
    ranges_d = torch.empty(max(nrangewords, 1), dtype=torch.int64, device=dev)

    def upload_and_broadcast():
        if rank == 0:
            words_d.copy_(words_h, non_blocking=True)
            if nrangewords:
                ranges_d[:nrangewords].copy_(ranges_h, non_blocking=True)
        if world > 1:
            dist.broadcast(words_d, 0)
            if nrangewords:
                dist.broadcast(ranges_d, 0)
        torch.cuda.synchronize()

    def open_iter():
        if alphabet == "dna":
            return sfx.Sfxiterator(sigma, n, None, world, twobit=words_d.data_ptr(), ranges=ranges_d.data_ptr(),
                                   numofwords=nwords, numofranges=nrangewords // 2, device=local, device_ptrs=True)
        return sfx.Sfxiterator(sigma, n, None, world, enc=words_d.data_ptr(), device=local, device_ptrs=True)

    upload_and_broadcast()
    it = open_iter()
    parts = it.parts
    if len(parts) != world:
        raise SystemExit(f"bucket table splits into {len(parts)} parts, {world} ranks")
    mincode, maxcode, width = parts[rank]
    stream = torch.cuda.ExternalStream(it.stream_handle(), device=dev)
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)   # > 126 MB L2
    info = sfx.CountInfo()

    class _DevArray:       # lets torch wrap the library's counter tables for the NCCL all-reduce
        def __init__(self, ptr, nwords):
            self.__cuda_array_interface__ = {"shape": (nwords,), "typestr": "<i4", "data": (ptr, False), "version": 3}

    shard = (n // world + 63) // 64 * 64
    pos_lo, pos_hi = min(n, rank * shard), (n if rank == world - 1 else min(n, (rank + 1) * shard))
    tables = {}
    # The all-reduce costs ~0.08 ms of launch + NCCL latency (measured at N = 2 on a 1 MB table: step 0.309 ms
    # replicated, 0.365 ms sharded), counting ~5 ps per position: sharding pays once the (1 - 1/N) n positions a
    # rank no longer counts are worth more than that.
    sharded_count = world > 1 and n * (1.0 - 1.0 / world) >= 24_000_000
    if os.environ.get("GTB200_BENCH_SHARDED_COUNT"):            # experiment knob: force either way
        sharded_count = world > 1 and os.environ["GTB200_BENCH_SHARDED_COUNT"] == "1"

    def count_step(ctx):
        """Counting phase: replicated at N = 1; position-sharded + one all-reduce of the sigma^pl tables at N > 1."""
        if not sharded_count:
            sfx._check(L.gtb200_count(ctx, None, None, None, None))   # tables stay on the device
            return
        ptr, nw = C.c_void_p(), C.c_uint64()
        sfx._check(L.gtb200_count_shard(ctx, pos_lo, pos_hi, 1 if rank == 0 else 0, C.byref(ptr), C.byref(nw)))
        key = (ptr.value, nw.value)
        if key not in tables:
            tables[key] = torch.as_tensor(_DevArray(ptr.value, nw.value), device=dev)
        dist.all_reduce(tables[key])                                  # ordered after the counting kernels on this stream
        sfx._check(L.gtb200_count_finish(ctx, None, None, None, None))

    def device_step():
        with torch.cuda.stream(stream):
            count_step(it.ctx)
        pinfo = it.sort_part_device(mincode, maxcode)
        if rank == world - 1:
            sfx._check(L.gtb200_tail(it.ctx, 0, it.specialcharacters + 1, None, 0))
        return pinfo

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        with torch.cuda.stream(stream):
            flush.zero_()
        device_step()
    def timed_pass(steps):
        """K steps, each bracketed by CUDA events on the library's stream, L2 flushed before every step."""
        evs = []
        last = None
        for _ in range(steps):
            with torch.cuda.stream(stream):
                flush.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(stream)
            last = device_step()
            b.record(stream)
            evs.append((a, b))
        barrier()
        return sum(a.elapsed_time(b) for a, b in evs), last

    # pass 1 (the reported value): un-instrumented.  Bracketing every kernel class with events costs ~25 % at
    # the 4.6 Mbp size (0.29 -> 0.36 ms per step), so the per-kernel times come from a second pass.
    it.profile(enable=False)
    barrier()
    sampler = ClockSampler(local) if rank == 0 else None
    t_wall0 = time.perf_counter()
    dev_ms, pinfo = timed_pass(args.steps)
    t_wall = time.perf_counter() - t_wall0
    # pass 2: the same K steps with CUDA events around every kernel class -> kernel_ms_per_step, roofline
    it.profile(enable=True)
    it.profile(reset=True)
    barrier()
    dev_ms_instr, _ = timed_pass(args.steps)
    clocks = sampler.stop() if sampler else None
    prof = it.profile()
    it_diag = it.part_diag()
    it.profile(enable=False)
    t = torch.tensor([dev_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms = float(t.item())
    ms_per_step = dev_ms / args.steps
    value = (n + 1) / (ms_per_step * 1e-3)

    # ---- e2e: host buffers in, host buffers out, through the Sfxiterator mirror ----
    e2e_steps = args.e2e_steps or max(2, min(args.steps, 5))
    suf_h = torch.empty(width + (it.specialcharacters + 1 if rank == world - 1 else 0), dtype=torch.int64).pin_memory()
    lcp_h = torch.empty(max(width, 1), dtype=torch.uint8).pin_memory()
    suf_np, lcp_np = suf_h.numpy().view(np.uint64), lcp_h.numpy()
    numcodes = sigma ** it.prefixlength
    lb_h = torch.empty(numcodes + 1, dtype=torch.int32).pin_memory()
    csc_h = torch.empty(sigma ** (it.prefixlength - 1), dtype=torch.int32).pin_memory()
    dist_h = torch.empty(max(1, int(it.countinfo.numofdistpfxidx)), dtype=torch.int32).pin_memory()
    it.delete()

    def e2e_step():
        upload_and_broadcast()                                  # pinned H2D (rank 0) + NVLink broadcast
        it2 = open_iter_raw()
        if not sharded_count:
            sfx._check(L.gtb200_count(it2, lb_h.data_ptr(), csc_h.data_ptr(), dist_h.data_ptr(), C.byref(info)))
        else:
            with torch.cuda.stream(torch.cuda.ExternalStream(int(L.gtb200_stream(it2)), device=dev)):
                ptr, nw = C.c_void_p(), C.c_uint64()
                sfx._check(L.gtb200_count_shard(it2, pos_lo, pos_hi, 1 if rank == 0 else 0, C.byref(ptr), C.byref(nw)))
                dist.all_reduce(torch.as_tensor(_DevArray(ptr.value, nw.value), device=dev))
            sfx._check(L.gtb200_count_finish(it2, lb_h.data_ptr(), csc_h.data_ptr(), dist_h.data_ptr(), C.byref(info)))
        pi = sfx.PartInfo()
        sfx._check(L.gtb200_sort_part(it2, mincode, maxcode, suf_np.ctypes.data, 0, lcp_np.ctypes.data, None, 0,
                                      C.byref(pi)))
        if rank == world - 1:
            sfx._check(L.gtb200_tail(it2, 0, int(info.specialcharacters) + 1, suf_np[width:].ctypes.data, 0))
        L.gtb200_close(it2)
        return pi

    def open_iter_raw():
        desc = sfx.SeqDesc()
        desc.numofchars, desc.totallength, desc.memory = sigma, n, 1
        if alphabet == "dna":
            desc.twobitencoding, desc.numofwords = words_d.data_ptr(), nwords
            desc.specialranges, desc.numofspecialranges = ranges_d.data_ptr(), nrangewords // 2
        else:
            desc.encoded = words_d.data_ptr()
        ctx = C.c_void_p()
        sfx._check(L.gtb200_open(C.byref(ctx), local, C.byref(desc), it.prefixlength))
        return ctx

    if args.skip_e2e:                       # profiling runs (ncu launch lists of the device steps only)
        e2e_steps, e2e_s = 0, float("nan")
    else:
        e2e_step()
        barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            e2e_step()
        barrier()
        e2e_s = (time.perf_counter() - t0) / e2e_steps
    t = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_s = float(t.item())
    h2d = (words_h.numel() * words_h.element_size() + ranges_h.numel() * 8) if rank == 0 else 0
    # the library ships 32-bit positions and widens them on the host when one GPU feeds one PCIe link (narrow transfer)
    narrow = os.environ.get("GTB200_NARROW", "1") != "0" and width >= (32 << 20)
    suf_bus_bytes = 4 if narrow else 8
    d2h = width * suf_bus_bytes + (suf_h.numel() - width) * 8 + width + (numcodes + 1) * 4 + csc_h.numel() * 4 \
        + int(it.countinfo.numofdistpfxidx) * 4
    tb = torch.tensor([h2d, d2h], dtype=torch.int64, device=dev)
    if world > 1:
        dist.all_reduce(tb)
    h2d, d2h = (int(x) for x in tb.tolist())

    if rank == 0:
        # ---- roofline of the dominant kernel class (rank 0's profile; all ranks do the same work) ----
        bits = 2 if alphabet == "dna" else 8
        text_bytes = n * bits / 8 + n / 8                       # packed text + special bitmap, read once
        m = width
        ncodes_part = maxcode - mincode + 1
        diag = it_diag                                          # what the library did for the last part (gtb200_part_diag)
        tile, levels, keyed = bool(diag["tile"]), int(diag["levels"]), bool(diag["keyed"])
        rec = 16 if keyed else 8                                # bytes per record of the tile path
        med = int(diag["medium_suffixes"])
        fb, rounds = int(pinfo.fallback_suffixes), int(pinfo.fallback_rounds)
        npad = int(diag["padded_suffixes"])
        algo = {   # algorithmic bytes per step and kernel class
            "count": text_bytes + 4 * numcodes,
            "scan": 8 * numcodes,
            "padded": 40 * npad + 12 * npad,                    # list built (key+pos in, radix passes aside) and placed
            # direct: text read + one 4-byte store; tile path: text read + records written once per level and
            # re-read once between levels
            "scatter": text_bytes + (4 * m if not tile else rec * m + (levels - 1) * 2 * rec * m),
            # direct: pos in, 8-byte text window, pos out, lcp out, .suf entry out; tile path reads a record instead
            # (keyed records bring the 8-byte window with them)
            "sort_lcp": ((4 + 8 if not tile else (16 if keyed else 8 + 8)) + 4 + 1 + 8) * max(m - med, 0),
            "sort_medium": (4 + 8 + 4 + 1 + 8) * med,
            # refinement: per round every open suffix passes ~11 radix passes of (8+4) bytes in and out
            "refine": fb * max(rounds, 1) * 11 * 24, "radix": fb * max(rounds, 1) * 11 * 24,
            "boundary": 8 * ncodes_part,
            "widen": 12 * m,
            "tail": 8 * (it.specialcharacters + 1),
            "part_init": 8 * ncodes_part,
        }
        kern = {k: v for k, v in prof.items() if v[1] > 0 and k != "prep"}
        total_ms = sum(v[0] for v in kern.values())
        dominant = max(kern, key=lambda k: kern[k][0])          # the class with the most device time, whatever it is
        d_ms, d_launches = kern[dominant]
        launches_per_step = max(1, d_launches // args.steps)
        per_launch_s = d_ms * 1e-3 / d_launches
        achieved = algo[dominant] / launches_per_step / per_launch_s / 1e9
        peak, peak_src = peak_hbm()
        traffic = ncu_traffic(args.workload, dominant) if world == 1 and SHRINK == 1 else None
        per_class = {k: {"ms": round(v[0] / args.steps, 4),
                         "frac": round(algo[k] / (v[0] / args.steps * 1e-3) / 1e9 / peak, 4) if v[0] > 0 and k in algo else None}
                     for k, v in kern.items()}
        line = {
            "metric": "suffixes/sec for SA+LCP", "value": value, "unit": "suffixes/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": args.scaling, "vs_baseline": None, "dtype": "u32 positions / u64 keys", "data": "synthetic",
            "config": {"workload": WORKLOADS[args.workload][1] + (f" x{world} (genome grows with N)" if world > 1 and args.scaling == "weak" else "")
                       + (f" SHRUNK 1/{SHRINK}" if SHRINK > 1 else ""),
                       "totallength": n, "prefixlength": it.prefixlength, "suffixes": n + 1,
                       "sharding": f"{world} bucket ranges (-parts split), text broadcast over NCCL, count "
                                   + ("position-sharded + all-reduce" if sharded_count else "replicated"),
                       "l2": "flushed (256 MiB write) before every timed step",
                       "passes": "value: K un-instrumented steps; kernel_ms_per_step/roofline: K more steps with CUDA events "
                                 "around every kernel class (that bracketing alone costs ~25 % at 4.6 Mbp)",
                       "fallback_suffixes": int(pinfo.fallback_suffixes),
                       "insertion": ("tile path: staged partition x%d (%d-byte records%s) + shared-memory tile sort"
                                     % (levels, rec, ", key travels with the record" if keyed else "")) if tile
                                    else "direct scatter + warp bucket sort",
                       "excluded_from_value": "gtb200_open's one-off device prep (special bitmap, range table, readmode "
                                              "transform): kernel class prep; e2e includes it"},
            "e2e": None if args.skip_e2e else
                   {"value": (n + 1) / e2e_s, "unit": "suffixes/s", "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": d2h, "ms_per_step": e2e_s * 1e3, "steps": e2e_steps,
                    "suf_entry_bytes_on_the_bus": suf_bus_bytes,
                    "host_result": "GtUword[suffixes] .suf entries + uint8 lcp + .bck tables in host memory"},
            "gpu_launches": int(sum(v[1] for v in kern.values())),
            "clocks": clocks,
            "roofline": {"bound": "hbm", "kernel": dominant, "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak,
                         "traffic": (traffic / launches_per_step) if traffic is not None else None,
                         "traffic_source": "profiles/ncu_traffic.json (ncu launch list of this command)" if traffic is not None else None,
                         "per_class": per_class,
                         "peak_source": peak_src,
                         "algorithmic_bytes_per_launch": algo[dominant] / launches_per_step,
                         "avg_launch_ms": per_launch_s * 1e3,
                         "share_of_step": d_ms / total_ms if total_ms else None,
                         "whole_path_frac": (n + 1) / world * (9.25 if alphabet == "dna" else 9.625)
                                            / (ms_per_step * 1e-3) / 1e9 / peak},
            "kernel_ms_per_step": {k: round(v[0] / args.steps, 5) for k, v in kern.items()},
            "kernel_events_pass_ms_per_step": dev_ms_instr / args.steps,
            "wall_s_timed_region": t_wall,
        }
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(enc, alphabet)
            full = full_size_reference(args.workload) if SHRINK == 1 else None
            if full:
                line["cpu_baseline"]["full_size_run"] = full
        print(json.dumps(line))
    if args.verify and world > 1:
        import hashlib
        itv = open_iter()
        with torch.cuda.stream(torch.cuda.ExternalStream(itv.stream_handle(), device=dev)):
            count_step(itv.ctx)
        pv = sfx.PartInfo()
        sufv = np.empty(width, dtype=np.uint64)
        lcpv = np.empty(width, dtype=np.uint8)
        sfx._check(L.gtb200_sort_part(itv.ctx, mincode, maxcode, sufv.ctypes.data, 0, lcpv.ctypes.data, None, 0, C.byref(pv)))
        mine = (int(pv.partoffset), width, hashlib.sha256(sufv.data).hexdigest(), hashlib.sha256(lcpv.data).hexdigest())
        itv.delete()
        allparts = [None] * world
        dist.all_gather_object(allparts, mine)
        if rank == 0:
            # single-context run of the whole genome on rank 0 (one part, prefix doubling available)
            if alphabet == "dna":
                itf = sfx.Sfxiterator(sigma, n, None, 1, twobit=words_d.data_ptr(), ranges=ranges_d.data_ptr(),
                                      numofwords=nwords, numofranges=nrangewords // 2, device=local, device_ptrs=True)
            else:
                itf = sfx.Sfxiterator(sigma, n, None, 1, enc=words_d.data_ptr(), device=local, device_ptrs=True)
            fsuf, flcp, _, _ = itf.next()
            itf.delete()
            ok = True
            for off, w, hs, hl in allparts:
                ok &= hashlib.sha256(fsuf[off:off + w].data).hexdigest() == hs
                ok &= hashlib.sha256(flcp[off:off + w].data).hexdigest() == hl
            ok &= sum(w for _, w, _, _ in allparts) == fsuf.shape[0]
            print("VERIFY multi-GPU parts vs single context:", "PASS" if ok else "FAIL", file=sys.stderr)
            if not ok:
                raise SystemExit("multi-GPU verification failed")
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
