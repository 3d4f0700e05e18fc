"""tests/cpp/reference_comparison_test.cpp -- a C++ program in the shape of the reference's on-card
test (host/test/RefernceComparisonTest/ReferenceComparisonTest.cpp:52-176), compiled against the
driver libraries and the oracle, run on generated files: two-way exact match of the 4-tuples, and
the abort path."""
import os
import subprocess

import pytest

import oracle_lib as O
from havac_b200 import build, synth

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = os.path.join(HERE, "golden")


def program():
    build.build_all()
    O.oracle()  # builds oracle/liboracle.so if it is missing
    src = os.path.join(HERE, "cpp", "reference_comparison_test.cpp")
    exe = os.path.join(HERE, "cpp", "reference_comparison_test")
    deps = [src, os.path.join(build.LIB, "libhavac_host.so"), os.path.join(O.ORACLE_DIR, "liboracle.so")]
    deps += [os.path.join(build.HOST, f) for f in os.listdir(build.HOST) if f.endswith(".hpp")]
    if not os.path.exists(exe) or os.path.getmtime(exe) < max(os.path.getmtime(d) for d in deps):
        subprocess.run(["g++", "-O2", "-std=c++17", src, "-I", build.HOST, "-I", build.INCLUDE, "-L", build.LIB, "-L", O.ORACLE_DIR,
                        "-lhavac_host", "-lhavac_cuda", "-loracle", f"-Wl,-rpath,{build.LIB}", f"-Wl,-rpath,{O.ORACLE_DIR}",
                        "-o", exe], check=True)
    return exe


def run(args):
    return subprocess.run(args, capture_output=True, text=True, timeout=600)


@pytest.mark.parametrize("name", ["two_models", "three_records"])
def test_golden_files(name):
    r = run([program(), "", os.path.join(GOLDEN, name + ".fa"), os.path.join(GOLDEN, name + ".hmm"), "0.02"])
    assert r.returncode == 0, r.stdout[-2000:]
    assert "found all" in r.stdout and "ERROR" not in r.stdout


def test_generated_pair_and_abort(tmp_path):
    models = synth.random_models(98, [300, 120, 210, 64])
    hmm = str(tmp_path / "m.hmm")
    synth.write_hmm(hmm, models)
    sym = synth.random_symbols(98, 2_000_000)
    synth.plant_homologs(98, sym, models, copies=12, substitution=0.1)
    fa = str(tmp_path / "s.fa")
    synth.write_fasta(fa, synth.split_records(sym, 5))
    r = run([program(), "", fa, hmm, "0.02"])
    assert r.returncode == 0, r.stdout[-2000:]
    hw = [l for l in r.stdout.splitlines() if l.startswith("found all")]
    assert len(hw) == 2 and int(hw[0].split()[2]) >= 12
    for p in ("0.1", "0.0001"):
        assert run([program(), "", fa, hmm, p]).returncode == 0
    r = run([program(), "", "abort", fa, hmm])
    assert r.returncode == 0 and "aborted. hw state is: " in r.stdout, r.stdout[-1000:]
