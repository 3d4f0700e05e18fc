"""bench.py's reference arm prints ONE JSON line with the keys the driver reads (CPU only: the
timed CPU search is replaced by a stub, the workload and the line's construction are the real ones)."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def test_reference_arm_line(monkeypatch, capsys):
    import bench

    seen = {}

    def fake_rate(packed, table, n_symbols, rows, threads, budget):
        seen.update(n=n_symbols, rows=rows, threads=threads, table_rows=table.shape[0])
        return 3.5, "reference", 2.0, 7

    monkeypatch.setattr(bench, "cpu_ssv_rate", fake_rate)
    args = argparse.Namespace(gpus=1, steps=2, warmup=1, impl="reference")
    bench.run_reference(args, rank=1)  # other ranks do nothing
    assert capsys.readouterr().out == ""
    bench.run_reference(args, rank=0)
    lines = [ln for ln in capsys.readouterr().out.splitlines() if ln.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == bench.METRIC and d["unit"] == bench.UNIT
    assert d["higher_is_better"] is True and d["n_gpus"] == 1 and d["steps"] == 2 and d["warmup"] == 1
    assert d["value"] == 3.5 and d["ms_per_step"] == 2000.0 and d["gpu_launches"] == 0
    assert d["config"]["workload"] == bench.workload_name(1, seen["table_rows"])
    cb = d["cpu_baseline"]
    assert cb["kind"] == "reference" and cb["cores"] == seen["threads"] and cb["value"] == d["value"] and cb["sample"]
    assert d["e2e"] == {"value": 3.5, "unit": bench.UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    # the sample is a slice of the workload: a bounded number of symbols per host thread, the first rows of the table
    assert seen["n"] == 250_000 * seen["threads"] and seen["rows"] == min(8000, seen["table_rows"])


def test_bench_table_is_the_reference_projection(monkeypatch):
    """bench.py's score table is what its label says: random HMMER3 models written to a .hmm file and projected
    to threshold 256 at p = 0.02 -- byte for byte what the UNMODIFIED reference objects (P7HmmReader +
    PhmmReprojection + PhmmPreprocessor, compiled into oracle/_ref) make of the same file."""
    import numpy as np
    import pytest

    import bench
    import oracle_lib as O
    from havac_b200 import synth

    if O.ref() is None:
        pytest.skip("oracle/_ref not built (no /root/reference here)")
    monkeypatch.setenv("HAVAC_BENCH_KEEP_HMM", "1")
    lengths = synth.model_lengths(bench.SEED, bench.N_MODELS, bench.LEN_LO, bench.LEN_HI)
    table, models = bench.build_table(lengths, 1000 + bench.SEED, "c2test")
    path = bench.workload_hmm_path("c2test")
    try:
        want, lens = O.ref_load_phmm(path, 0.02)
    finally:
        os.unlink(path)
    assert lens.tolist() == [int(x) for x in lengths]
    assert np.array_equal(table.reshape(-1), want)
    # and build_workload uses exactly this table
    monkeypatch.delenv("HAVAC_BENCH_KEEP_HMM")
    _, table2, n, _ = bench.build_workload(1)
    assert n == bench.SEQ_PER_GPU and np.array_equal(table2, table)
