"""world_size-2 (and 3) CPU tests of the multi-GPU host logic over gloo: every rank derives the same
shard plan from rank 0's bucket table, shards tile the suffix array without gaps, and the gathered
statistics equal the single-part values of the oracle."""
import os
import socket

import numpy as np
import pytest
import torch.distributed as dist
import torch.multiprocessing as mp

from genometools_b200 import shard, synth
from oracle import oracle


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        enc = synth.multi_dna_with_nruns(60_000, nseq=5, seed=77)
        o = oracle.build(enc, 4)                      # every rank can afford the oracle at this size
        lb0 = o["leftborder"] if rank == 0 else None
        plan, lb = shard.agree_on_plan(lb0, o["prefixlength"])
        assert np.array_equal(lb, o["leftborder"])
        lo, hi, width, offset = plan[rank]
        # what a rank would get back from gtb200_sort_part for its range, taken from the oracle
        lcp = o["lcp_array"][offset:offset + width].astype(np.float64)
        suf = o["suf_array"][offset:offset + width]
        # lcptabsum counts only suffixes with >= pl real symbols: mask via distance to next special
        sp = np.concatenate([np.flatnonzero(enc >= 254), [enc.shape[0]]])
        dist_to_sp = sp[np.searchsorted(sp, suf)] - suf
        mysum = float(lcp[dist_to_sp >= o["prefixlength"]].sum())
        hit = np.flatnonzero(suf == 0)
        mylongest = offset + int(hit[0]) if hit.size else -1
        tot = shard.gather_shard_stats(width, mysum, int(lcp.max()) if width else 0,
                                       int((lcp >= 255).sum()), mylongest)
        q.put((rank, plan, tot, o["nonspecial"], o["averagelcp"], o["maxbranchdepth"], o["longest"], enc.shape[0]))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_shard_plan_and_stats_over_gloo(world):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    plans = [r[1] for r in sorted(res)]
    assert all(p == plans[0] for p in plans), "ranks disagree on the shard plan"
    plan = plans[0]
    _, _, tot, nonspecial, averagelcp, maxdepth, longest, n = res[0]
    assert plan[0][0] == 0 and plan[0][3] == 0
    for (lo0, hi0, w0, off0), (lo1, hi1, w1, off1) in zip(plan[:-1], plan[1:]):
        assert lo1 == hi0 + 1 and off1 == off0 + w0
    assert sum(p[2] for p in plan) == nonspecial == tot[0]
    assert "%.2f" % (tot[1] / (n + 1)) == averagelcp
    assert tot[2] == maxdepth
    first_is_special = synth.multi_dna_with_nruns(60_000, nseq=5, seed=77)[0] >= 254
    # suffix 0 sits in the tail (owned by the last rank's gtb200_tail call) when position 0 is a special
    assert tot[4] == (-1 if first_is_special else longest)
