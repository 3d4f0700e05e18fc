"""tools/window_check.py: the logic of the reference's HMMER cross-check
(test/hmmerSsvRef/hmmerSsvRef.cpp:166-231, 379-413, 504-586) fed with a window list made here from
the oracle's hits, since no nhmmer is installed: windows that end on an oracle hit and cover only
part of its scoring stretch (so the walk-back is exercised), plus decoys on random diagonals."""
import os
import sys

import numpy as np
import pytest

import oracle_lib as O
from havac_b200 import havac, synth

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))


def test_traceback_walks_back_and_forward():
    import window_check as W
    table = np.full((40, 4), -20, dtype=np.int8)
    table[:, 0] = 10
    sym = np.zeros(100, dtype=np.uint8)  # all A: +10 per step on any diagonal
    # a 6-step window scores 70; the walk back over rows 0..13 adds 140, the walk forward reaches 256
    assert W.traceback_score(sym, table, 30, 36, 20) >= 256
    sym[31] = 1  # a mismatch inside the window does not stop the walk
    assert W.traceback_score(sym, table, 30, 36, 20) >= 256
    sym[:] = 1
    assert 0 <= W.traceback_score(sym, table, 30, 36, 20) < 256
    assert W.traceback_score(sym, table, 30, 36, 3) == -1  # the window would start above row 0


def test_window_list_against_a_gpu_run(tmp_path):
    import window_check as W
    models = synth.random_models(97, [200, 310, 150])
    hmm = str(tmp_path / "m.hmm")
    synth.write_hmm(hmm, models)
    sym = synth.random_symbols(97, 360_000)
    synth.plant_homologs(97, sym, models, copies=8, substitution=0.08)
    fa = str(tmp_path / "s.fa")
    synth.write_fasta(fa, synth.split_records(sym, 2))

    packed, ends = havac.pack_fasta(fa, seed=1)
    table, lens = havac.project_phmm(hmm, 0.02)
    names = havac.model_names(hmm)
    pos, row = O.soft_ssv(synth.unpack_2bit(packed), table)
    first_row = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    starts = np.concatenate([[0], ends[:-1]]).astype(np.int64)
    g = synth.rng(97, 9)
    lines = ["sequenceId\twindowStart\twindowEnd\twindowLength\thmmPosition\taccession"]
    made = 0
    for p, r in zip(pos.tolist(), row.tolist()):
        rec = int(np.searchsorted(ends, p, side="left"))
        if rec >= len(ends):
            continue
        model = int(np.searchsorted(first_row, r, side="right") - 1)
        we, hp = p - int(starts[rec]), r - int(first_row[model])
        length = min(we, hp, int(g.integers(5, 80)))
        if length < 1:
            continue
        lines.append(f"{rec}\t{we - length + 1}\t{we + 1}\t{length}\t{hp - 1}\t{names[model][0]}")
        made += 1
    assert made >= 8
    decoys = 40
    for _ in range(decoys):
        rec, model = int(g.integers(0, len(ends))), int(g.integers(0, len(lens)))
        hp = int(g.integers(40, lens[model] - 1))
        we = int(g.integers(100, int(ends[rec] - starts[rec]) - 100))
        lines.append(f"{rec}\t{we - 30 + 1}\t{we + 1}\t30\t{hp - 1}\t{names[model][0]}")
    lines.append("not a window line")
    tsv = tmp_path / "windows.tsv"
    tsv.write_text("\n".join(lines) + "\n")

    windows = W.read_windows(str(tsv))
    assert len(windows) == made + decoys
    report = W.check(fa, hmm, windows, 0.02)
    assert report["verified_but_missed"] == []
    # the walk is the reference's heuristic, not the recurrence: a window cut out of a stretch that
    # had been reset by an earlier hit on the same diagonal can stay short of 256
    assert report["verified"] >= 0.95 * made
    assert report["verified_and_reported"] == report["verified"]
    assert report["below_threshold"] >= decoys - 3 and report["off_model"] == 0
    assert report["verified"] + report["below_threshold"] == made + decoys
    assert report["havac_hits"] == int((pos < ends[-1] + 1).sum())
