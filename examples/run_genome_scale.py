"""BASELINE configs[2] from FILES through the C++ user program: a 3 Gbp FASTA (24 records sized like
hg38's chromosomes) against a 3,000-model HMMER3 database, every GPU of the box ("cuda:all").
Whole-program seconds include device open, both file loads, the search and hit resolution.
Appends one JSON line to gpurun_out/genome_scale.jsonl.   usage: run_genome_scale.py [Gbp] [models]"""
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "examples"))
from havac_b200 import synth  # noqa: E402
import run_benchmark  # noqa: E402

HG38_MBP = [248, 242, 198, 190, 182, 171, 159, 145, 138, 134, 135, 133, 114, 107, 102, 90, 83, 80, 59, 64, 47, 51, 156, 57]


def write_big_fasta(path, total, seed):
    """80-column FASTA of i.i.d. uniform DNA, written in 64 M-symbol pieces."""
    g = synth.rng(seed, 11)
    letters = np.frombuffer(b"ACGT", dtype=np.uint8)
    sizes = [int(total * m / sum(HG38_MBP)) for m in HG38_MBP]
    sizes[-1] += total - sum(sizes)
    with open(path, "wb") as fh:
        for k, n in enumerate(sizes):
            fh.write(b">chr%d synthetic\n" % (k + 1))
            left = n
            while left > 0:
                m = min(left, 64_000_000 // 80 * 80)
                sym = letters[g.integers(0, 4, size=m, dtype=np.uint8)]
                full = m // 80 * 80
                lines = np.empty((full // 80, 81), dtype=np.uint8)
                lines[:, :80] = sym[:full].reshape(-1, 80)
                lines[:, 80] = 10
                fh.write(lines.tobytes())
                if m > full:
                    fh.write(sym[full:].tobytes() + b"\n")
                left -= m
    return sizes


def main():
    gbp = float(sys.argv[1]) if len(sys.argv) > 1 else 3.0
    n_models = int(sys.argv[2]) if len(sys.argv) > 2 else 3000
    exe = run_benchmark.build_program()
    with tempfile.TemporaryDirectory() as tmp:
        fa, hmm = os.path.join(tmp, "genome.fa"), os.path.join(tmp, "db.hmm")
        t0 = time.time()
        write_big_fasta(fa, int(gbp * 1e9), 3)
        lens = synth.model_lengths(3, n_models, 100, 700)
        synth.write_hmm(hmm, synth.random_models(3, [int(x) for x in lens]))
        t_files = time.time() - t0
        spec = "cuda:all;hitcap=48000000+"
        t0 = time.time()
        out = subprocess.run([exe, spec, fa, hmm, "0.02"], check=True, capture_output=True, text=True).stdout
        wall = time.time() - t0
        line = json.loads([l for l in out.splitlines() if l.startswith("{")][-1])
        line.update({"case": f"{gbp} Gbp FASTA x {n_models} models from files, {spec}", "fasta_bytes": os.path.getsize(fa),
                     "hmm_bytes": os.path.getsize(hmm), "make_files_s": round(t_files, 1), "process_wall_s": round(wall, 2)})
        for k in ("fasta", "hmm"):
            line[k] = os.path.basename(line[k])
        print(json.dumps(line), flush=True)
        d = os.path.join(ROOT, "gpurun_out")
        if os.path.isdir(d):
            with open(os.path.join(d, "genome_scale.jsonl"), "a") as fh:
                fh.write(json.dumps(line) + "\n")


if __name__ == "__main__":
    main()
