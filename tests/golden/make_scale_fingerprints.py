"""Generates tests/golden/scale_fingerprints.json with the UNMODIFIED reference (oracle/_ref/gt_ref, built by
oracle/Makefile from /root/reference) at BASELINE.json's FULL sizes.  Run in the build container only (CPU):

    python tests/golden/make_scale_fingerprints.py chr1 protein human      # 80 s + ~10 min + ~20 min of one core

For every workload it regenerates the synthetic input from its seed (bench.make_workload), writes it as FASTA,
runs `gt suffixerator -dna|-protein -suf -lcp -bck -tis`, and records size + SHA-256 of .suf/.lcp/.llv/.bck and the
.prj statistics.  tests/test_gpu_scale.py rebuilds the same index through the C ABI on the GPU box and compares
the hashes: byte identity at full size without shipping 28 GB of golden files.
"""
import hashlib
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "scale_fingerprints.json")


def file_sha(path, chunk=1 << 26):
    h = hashlib.sha256()
    with open(path, "rb") as fh:
        while True:
            b = fh.read(chunk)
            if not b:
                break
            h.update(b)
    return h.hexdigest()


def write_input(name, fasta):
    """Child process: synthesise + write FASTA, so that its memory is gone before the reference starts."""
    from bench import make_workload
    from tests_scale_common import fast_write_fasta
    enc, alphabet = make_workload(name, 1)
    fast_write_fasta(fasta, enc, alphabet)
    print(json.dumps({"alphabet": alphabet, "totallength": int(enc.shape[0]),
                      "specialcharacters": int((enc >= 254).sum()),
                      "enc_sha256": hashlib.sha256(enc.data).hexdigest()}))


def run_one(name, oracle):
    with tempfile.TemporaryDirectory(dir=os.environ.get("GT_SCALE_TMP", "/tmp")) as tmp:
        fa = os.path.join(tmp, "w.fna")
        t0 = time.time()
        meta = json.loads(subprocess.check_output([sys.executable, os.path.abspath(__file__), "--write-input", name, fa],
                                                  text=True).splitlines()[-1])
        print(f"{name}: n={meta['totallength']} fasta {time.time()-t0:.0f}s", flush=True)
        cmd = [oracle.GT_REF, "suffixerator", "-" + meta["alphabet"], "-suf", "-lcp", "-bck", "-tis", "-v",
               "-indexname", os.path.join(tmp, "ref"), "-db", fa]
        t0 = time.time()
        p = subprocess.run(cmd, capture_output=True, text=True, env=dict(os.environ, GT_ENV_OPTIONS="-showtime"))
        wall = time.time() - t0
        if p.returncode != 0:
            raise SystemExit(f"reference failed on {name}: {p.stderr[-2000:]}")
        times = {}
        for line in p.stdout.splitlines():
            if line.startswith("# TIME "):
                k, _, v = line[7:].rpartition(" ")
                try:
                    times[k] = float(v)
                except ValueError:
                    pass
        prj = oracle.parse_prj(open(os.path.join(tmp, "ref.prj")).read())
        rec = dict(meta)
        rec["files"] = {k: {"bytes": os.path.getsize(os.path.join(tmp, "ref." + k)),
                            "sha256": file_sha(os.path.join(tmp, "ref." + k))} for k in ("suf", "lcp", "llv", "bck")}
        rec["prj"] = {k: prj[k] for k in ("totallength", "specialcharacters", "longest", "prefixlength", "largelcpvalues",
                                          "averagelcp", "maxbranchdepth") if k in prj}
        rec["reference"] = {"wall_s": round(wall, 1), "times_s": times, "threads": 1,
                            "command": "gt suffixerator -%s -suf -lcp -bck -tis -v" % meta["alphabet"]}
        return rec


def main():
    names = sys.argv[1:] or ["chr1", "protein", "human"]
    if names[0] == "--write-input":
        write_input(names[1], names[2])
        return
    from oracle import oracle
    for name in names:
        rec = run_one(name, oracle)
        db = json.load(open(OUT)) if os.path.exists(OUT) else {}
        db[name] = rec
        print(json.dumps({name: rec}), flush=True)
        with open(OUT, "w") as fh:
            json.dump(db, fh, indent=1, sort_keys=True)
            fh.write("\n")


if __name__ == "__main__":
    main()
