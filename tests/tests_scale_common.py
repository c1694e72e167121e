"""Shared by the scale-check scripts (not collected by pytest)."""
import numpy as np

from genometools_b200 import encseq


def fast_write_fasta(path, enc, alphabet, width=80):
    seqs = encseq.decode_to_fasta_bytes(enc, alphabet)
    with open(path, "wb") as fh:
        for k, s in enumerate(seqs):
            fh.write(b">seq%d\n" % k)
            a = np.frombuffer(s, dtype=np.uint8)
            full = (a.shape[0] // width) * width
            step = width * (1 << 20)
            for lo in range(0, full, step):
                hi = min(full, lo + step)
                body = np.empty(((hi - lo) // width, width + 1), dtype=np.uint8)
                body[:, :width] = a[lo:hi].reshape(-1, width)
                body[:, width] = 10
                fh.write(body.tobytes())
            if a.shape[0] > full:
                fh.write(a[full:].tobytes() + b"\n")
