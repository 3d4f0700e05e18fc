"""Where the time of getHitsFromFinishedRun goes at tens of millions of hits: a hit-dense run
through the C++ user program with HAVAC_TRACE=1 (stage timings on stderr).  usage: [Mbp] [p]"""
import os
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "examples"))
from havac_b200 import synth  # noqa: E402
import run_benchmark  # noqa: E402

mbp = int(sys.argv[1]) if len(sys.argv) > 1 else 100
p = sys.argv[2] if len(sys.argv) > 2 else "0.9"
exe = run_benchmark.build_program()
with tempfile.TemporaryDirectory() as tmp:
    sym = synth.random_symbols(2, mbp * 1_000_000)
    models = synth.random_models(2, synth.model_lengths(2, 100, 100, 600))
    fa, hmm = os.path.join(tmp, "s.fa"), os.path.join(tmp, "m.hmm")
    synth.write_fasta(fa, synth.split_records(sym, 6))
    synth.write_hmm(hmm, models)
    env = dict(os.environ, HAVAC_TRACE="1")
    r = subprocess.run([exe, ";hitcap=64000000+", fa, hmm, p], env=env, capture_output=True, text=True)
    print(r.stdout[-700:])
    print("\n".join(l for l in r.stderr.splitlines() if "trace" in l))
