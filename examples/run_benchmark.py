"""Builds examples/havac_benchmark.cpp against the in-tree driver libraries and runs it on
synthetic files: (a) BASELINE configs[1] from files (100 Mbp FASTA x 100 models), and (b) the
shape of the reference's published end-to-end series (benchmark/runtime_table.py: one strand of
chr22, 50.8 Mbp, against model databases of 1 k ... 150 k positions).
Whole-program seconds include device open, both file loads, the search and hit resolution.
Writes one JSON line per run to gpurun_out/user_program.jsonl (when that directory exists)."""
import json
import os
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from havac_b200 import build, synth  # noqa: E402

PUBLISHED = {1007: 6.06, 5055: 6.31, 10122: 6.766, 20039: 6.88, 30007: 7.41, 50120: 8.339, 60156: 8.88,
             70107: 9.38, 80042: 9.91, 90003: 10.86, 100048: 11.61, 150043: 14.16}  # HAVAC on the U50, seconds


def build_program():
    build.build_all()
    exe = os.path.join(ROOT, "examples", "havac_benchmark")
    src = exe + ".cpp"
    # the program includes the driver's C++ headers (class layouts): rebuild when any of them, or a library, is newer
    deps = [src, os.path.join(build.LIB, "libhavac_host.so"), os.path.join(build.LIB, "libhavac_cuda.so")]
    deps += [os.path.join(build.HOST, f) for f in os.listdir(build.HOST) if f.endswith((".hpp", ".h"))]
    if not os.path.exists(exe) or os.path.getmtime(exe) < max(os.path.getmtime(d) for d in deps):
        subprocess.run(["g++", "-O2", "-std=c++17", src, "-I", build.HOST, "-I", build.INCLUDE, "-L", build.LIB,
                        "-lhavac_host", "-lhavac_cuda", f"-Wl,-rpath,{build.LIB}", "-o", exe], check=True)
    return exe


def models_with_total(seed, total, lo=60, hi=400):
    g = synth.rng(seed, 31)
    lens = []
    while sum(lens) < total:
        lens.append(int(g.integers(lo, hi + 1)))
    lens[-1] -= sum(lens) - total
    if lens[-1] < 10:
        lens[-2] += lens.pop()
    return synth.random_models(seed, lens)


def run(exe, fasta, hmm, extra, both=False):
    out = subprocess.run([exe, "", fasta, hmm, "0.02"] + (["both"] if both else []), check=True, capture_output=True,
                         text=True).stdout
    line = json.loads([l for l in out.splitlines() if l.startswith("{")][-1])
    line.update(extra)
    for k in ("fasta", "hmm"):
        line[k] = os.path.basename(line[k])
    print(json.dumps(line), flush=True)
    d = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(d):
        with open(os.path.join(d, "user_program.jsonl"), "a") as fh:
            fh.write(json.dumps(line) + "\n")


def main():
    exe = build_program()
    if "--build-only" in sys.argv:
        return
    with tempfile.TemporaryDirectory() as tmp:
        # (a) configs[1] through files
        sym = synth.random_symbols(2, 100_000_000)
        models = synth.random_models(2, synth.model_lengths(2, 100, 100, 600))
        synth.plant_homologs(2, sym, models, copies=10)
        fa, hmm = os.path.join(tmp, "c2.fa"), os.path.join(tmp, "c2.hmm")
        synth.write_fasta(fa, [sym])
        synth.write_hmm(hmm, models)
        run(exe, fa, hmm, {"case": "configs[1] from files: 100 Mbp x 100 models"})
        run(exe, fa, hmm, {"case": "configs[1] from files, both strands (reverse complement made on the device)"}, both=True)
        # (a') configs[0], the reference's own CPU-runnable case: 1 Mbp x one 200-position model; run_us is the latency
        # of Havac::runHardwareClient (launch, wait), get_hits_us that of getHitsFromFinishedRun
        sym = synth.random_symbols(1, 1_000_000)
        models = synth.random_models(1, [200])
        synth.plant_homologs(1, sym, models, copies=10)
        fa1, hmm1 = os.path.join(tmp, "c1.fa"), os.path.join(tmp, "c1.hmm")
        synth.write_fasta(fa1, [sym])
        synth.write_hmm(hmm1, models)
        run(exe, fa1, hmm1, {"case": "configs[0] from files: 1 Mbp x 1 model of 200 positions"})
        # (b) the published series' shape
        sym = synth.random_symbols(22, 50_800_000)
        fa = os.path.join(tmp, "chr22_like.fa")
        synth.write_fasta(fa, [sym])
        for total, published in PUBLISHED.items():
            models = models_with_total(total, total)
            hmm = os.path.join(tmp, f"db{total}.hmm")
            synth.write_hmm(hmm, models)
            run(exe, fa, hmm, {"case": "chr22-sized strand x model database", "model_positions": total,
                               "published_u50_total_s": published})


if __name__ == "__main__":
    main()
