"""Device-side score-table projection (havac_cuda_project_phmm) against the
reference's p7HmmProjectForThreshold256 (PhmmReprojection/PhmmReprojection.cpp:109-145): the int8
table the device builds from the models' negative-log emissions is read back and must equal, byte
for byte, the committed golden tables (made by the reference objects), the reference itself where
oracle/_ref is present, and the host projection (which the CPU suite pins to the reference)."""
import os
import time

import numpy as np
import pytest

import oracle_lib as O
from havac_b200 import havac, synth
from havac_b200.client import HavacHwClient

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = os.path.join(HERE, "golden")
LOG2_E = np.float32(1.44269504089)


def device_table(hmm, p):
    h = havac.Havac(0, p)
    try:
        h.loadPhmm(str(hmm))
        return h.devicePhmmScores()
    finally:
        h.close()


def restated(emissions, lengths, scales):
    """float32 alpha - (x * beta), C round() (half away from zero), saturate: PhmmReprojection.cpp:117-144."""
    out = np.zeros(emissions.shape, dtype=np.int8)
    at = 0
    with np.errstate(invalid="ignore", over="ignore"):
        for n, s in zip(lengths, scales):
            s = np.float32(s)
            alpha, beta = np.float32(2) * s, LOG2_E * s
            x = emissions[at:at + n].astype(np.float32)
            v = (alpha - (x * beta).astype(np.float32)).astype(np.float32).astype(np.float64)
            v = np.trunc(v + np.copysign(0.5, v))  # exact in double for a float32 argument
            out[at:at + n] = np.clip(v, -128, 127).astype(np.int8)
            at += n
    return out


@pytest.mark.parametrize("name", ["two_models", "three_records"])
def test_device_table_matches_golden(name):
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    hmm = os.path.join(GOLDEN, name + ".hmm")
    if not os.path.exists(hmm):
        pytest.skip("no .hmm beside this golden set")
    tab = device_table(hmm, float(z["p_value"]))
    assert np.array_equal(tab.reshape(-1), z["table"].reshape(-1))


@pytest.mark.parametrize("p", [0.02, 1e-4, 0.5, 1e-10])
def test_device_table_vs_host_projection_and_reference(tmp_path, p):
    lens = [1, 2, 7, 300, 64, 2048, 2049, 5, 1200]
    models = synth.random_models(71, lens)
    hmm = str(tmp_path / "m.hmm")
    synth.write_hmm(hmm, models)
    tab = device_table(hmm, p)
    host, hl = havac.project_phmm(hmm, p)
    assert list(hl) == lens and tab.shape == host.shape
    assert np.array_equal(tab, host)
    if O.ref() is not None:
        rt, rl = O.ref_load_phmm(hmm, p)
        assert np.array_equal(tab.reshape(-1), rt)


def test_rounding_ties_saturation_and_infinite_emissions():
    """Straight through the C-ABI with chosen scale factors: emissions placed so that
    alpha - x*beta lands on and next to every half-integer in range, far outside the int8 range,
    and at +inf (HMMER's '*')."""
    g = synth.rng(72, 1)
    scales = np.array([0.5, 1.0, 3.75, 0.0123, 40.0, 1.0 / 1.44269504089], dtype=np.float32)
    per = 4096
    rows = []
    for s in scales:
        alpha, beta = np.float32(2) * s, LOG2_E * s
        want = np.concatenate([np.arange(-140, 140, 0.5), g.uniform(-200, 200, size=per * 4 - 560 - 16)]).astype(np.float32)
        x = ((alpha - want) / beta).astype(np.float32)
        # neighbours one ulp either side so that the product's rounding is exercised
        x[1::3] = np.nextafter(x[1::3], np.float32(np.inf))
        x[2::3] = np.nextafter(x[2::3], np.float32(-np.inf))
        special = np.array([np.inf, 0.0, -0.0, 1e-38, 1e38, 3.4e38, -1e38, 88.0, 1e-45, 0.5, 1.0, 2.0, 4.0, 8.0, 16.0, 32.0], dtype=np.float32)
        rows.append(np.concatenate([x, special]).reshape(per, 4))
    e = np.concatenate(rows)
    lens = np.full(len(scales), per, dtype=np.uint32)
    c = HavacHwClient("", 0)
    try:
        c.projectPhmm(e, lens, scales)
        tab = c.readPhmm()
    finally:
        c.close()
    want = restated(e, lens, scales)
    assert np.array_equal(tab, want)
    assert (tab == 127).any() and (tab == -128).any()


def test_projected_table_runs_like_a_written_one(tmp_path):
    """Same hits whether the table is projected on the device or written as int8 from the host."""
    models = synth.random_models(73, [150, 333, 90])
    hmm = str(tmp_path / "m.hmm")
    synth.write_hmm(hmm, models)
    sym = synth.random_symbols(73, 500_000)
    synth.plant_homologs(73, sym, models, copies=9, substitution=0.1)
    fa = str(tmp_path / "s.fa")
    synth.write_fasta(fa, synth.split_records(sym, 2))

    def hits(host_side, monkey):
        if host_side:
            monkey.setenv("HAVAC_HOST_PROJECT", "1")
        else:
            monkey.delenv("HAVAC_HOST_PROJECT", raising=False)
        h = havac.Havac(0, 0.02)
        try:
            havac.srand(3)
            h.loadPhmm(hmm)
            h.loadSequence(fa)
            h.runHardwareClient()
            return sorted(h.getHitsFromFinishedRun())
        finally:
            h.close()

    with pytest.MonkeyPatch.context() as mp:
        dev, host = hits(False, mp), hits(True, mp)
    assert len(dev) >= 9 and dev == host


def test_project_errors():
    c = HavacHwClient("", 0)
    try:
        e = np.zeros((10, 4), dtype=np.float32)
        with pytest.raises(Exception):  # lengths do not add up
            c.projectPhmm(e, [4, 5], [1.0, 1.0])
        c.projectPhmm(e, [4, 6], [1.0, 1.0])
        assert c.readPhmm().shape == (10, 4) and (c.readPhmm() == 2).all()  # x = 0: alpha = 2
    finally:
        c.close()


def test_projection_at_the_largest_table_shape():
    """1.19 M rows (BASELINE config 3's 3,000 models): device projection vs the restatement, timed."""
    lens = synth.model_lengths(3, 3000, 100, 700).astype(np.uint32)
    rows = int(lens.sum())
    g = synth.rng(74, 1)
    e = g.exponential(1.6, size=(rows, 4)).astype(np.float32)
    scales = g.uniform(0.8, 3.0, size=len(lens)).astype(np.float32)
    c = HavacHwClient("", 0)
    try:
        c.projectPhmm(e[:1000], [1000], scales[:1])  # warm
        t0 = time.perf_counter()
        c.projectPhmm(e, lens, scales)
        dt = time.perf_counter() - t0
        tab = c.readPhmm()
    finally:
        c.close()
    t0 = time.perf_counter()
    want = restated(e, lens, scales)
    dt_np = time.perf_counter() - t0
    assert np.array_equal(tab, want)
    print(f"project {rows} rows: device {dt * 1e3:.2f} ms, numpy restatement {dt_np * 1e3:.1f} ms")
