"""GPU test of gtb200_leftborder_partialsums_noalloc, the one entry point that never allocates
device memory (for hosts whose own pool owns the whole GPU, or that cannot tolerate the hidden
device synchronisation of cudaMalloc/cudaFree).  It must give exactly the bucket table the regular
path (gtb200_count) gives - integers, so the comparison is bit-exact like every other parity test
here - on two calls in a row that share the library's fixed scratch array, and it must refuse
tables larger than that array.  The file name keeps it last in collection order; it does not
depend on any other test having run (or not run) in the process."""
import ctypes as C

import numpy as np
import pytest

from genometools_b200 import encseq, sfx, synth

pytestmark = pytest.mark.gpu


def _table_from_regular_path(enc, pl):
    it = sfx.Sfxiterator(4, enc.shape[0], pl, 1, twobit=encseq.twobit_pack(enc), ranges=encseq.special_ranges(enc))
    lb = it.leftborder.copy()
    it.delete()
    return lb


def _noalloc(counts_t, out_t, numcodes, stream_handle):
    L = sfx.load_library()
    err = C.create_string_buffer(256)
    rc = L.gtb200_leftborder_partialsums_noalloc(counts_t.data_ptr(), out_t.data_ptr(), numcodes, stream_handle,
                                                 err, len(err))
    return rc, err.value.decode()


def test_noalloc_prefix_sum_matches_regular_path_twice_in_a_row():
    import torch

    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    stream = torch.cuda.current_stream(dev)
    # first call: a large table (4^9 buckets, 65 tiles of partial sums); second call right after it:
    # a small one (4^5 buckets, one tile), so anything stale in the shared scratch would show
    cases = [(synth.multi_dna_with_nruns(2_000_000, nseq=6, seed=71), 9), (synth.random_dna(30_000, seed=72), 5)]
    wants = [_table_from_regular_path(enc, pl) for enc, pl in cases]
    # the test harness may allocate; the entry point under test does not
    ins = [torch.from_numpy(np.diff(w.astype(np.int64)).astype(np.int32)).to(dev) for w in wants]
    outs = [torch.full((w.shape[0],), -1, dtype=torch.int32, device=dev) for w in wants]
    torch.cuda.synchronize()
    for (enc, pl), cin, cout in zip(cases, ins, outs):      # both calls back to back on one stream
        rc, msg = _noalloc(cin, cout, 4 ** pl, stream.cuda_stream)
        assert rc == 0, msg
    torch.cuda.synchronize()
    for want, cout in zip(wants, outs):
        got = cout.cpu().numpy().view(np.uint32)
        assert np.array_equal(got, want), "no-alloc prefix sum differs from gtb200_count's leftborder"
    # in place (output buffer = input buffer with one spare entry) and a repeat of the first size
    buf = torch.zeros(wants[0].shape[0], dtype=torch.int32, device=dev)
    buf[:-1] = ins[0]
    rc, msg = _noalloc(buf, buf, 4 ** cases[0][1], stream.cuda_stream)
    assert rc == 0, msg
    torch.cuda.synchronize()
    assert np.array_equal(buf.cpu().numpy().view(np.uint32), wants[0])


def test_noalloc_rejects_tables_beyond_its_fixed_scratch():
    import torch

    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    t = torch.zeros(16, dtype=torch.int32, device=dev)
    rc, msg = _noalloc(t, t, 4 ** 13 + 1, 0)     # rejected before anything is read or launched
    assert rc == -1 and "exceed" in msg and "67108864" in msg
    rc, msg = _noalloc(t, t, 0, 0)
    assert rc == -1


def test_noalloc_on_a_second_gpu():
    """The fixed scratch array exists once per GPU and its address is resolved per device: the same call on device 1
    (after device 0 has been used) must give device 1's table, not touch device 0's array."""
    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs a second GPU")
    enc = synth.multi_dna_with_nruns(1_000_000, nseq=4, seed=73)
    want = _table_from_regular_path(enc, 8)
    for d in (0, 1, 0):
        dev = torch.device("cuda", d)
        torch.cuda.set_device(dev)
        cin = torch.from_numpy(np.diff(want.astype(np.int64)).astype(np.int32)).to(dev)
        cout = torch.full((want.shape[0],), -1, dtype=torch.int32, device=dev)
        torch.cuda.synchronize(dev)
        rc, msg = _noalloc(cin, cout, 4 ** 8, torch.cuda.current_stream(dev).cuda_stream)
        assert rc == 0, msg
        torch.cuda.synchronize(dev)
        assert np.array_equal(cout.cpu().numpy().view(np.uint32), want), f"device {d}"
    torch.cuda.set_device(0)
