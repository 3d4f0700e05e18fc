"""Cross-check of a HAVAC run against a pre-dumped list of nhmmer SSV windows.

The reference's test/hmmerSsvRef/hmmerSsvRef.cpp does this against a live HMMER install: it reads
a tab-separated window list (`readHmmerHits`, :379-413: one header line, then
`sequenceId  windowStart  windowEnd  windowLength  hmmPosition  accession`, 1-based sequence
coordinates), re-scores every window along its diagonal with the projected int8 scores, walking
back from the window start and forward from its end if the window alone does not reach 256
(`attemptTraceback`, :504-586), and counts the windows the threshold-256 projection can verify
(`checkHmmerHitsAgainstRefSsv`, :166-231). HMMER is not on this box, so the window list comes from a
file made elsewhere; this tool runs the search on the GPU through the Havac class, re-scores the
windows on the CPU with the same walk, and reports

  * windows whose diagonal reaches 256 (the projection agrees with nhmmer's filter),
  * of those, windows for which HAVAC reported a hit on that diagonal (it must: a miss is a bug),
  * windows that stay below 256 (nhmmer's 8-bit filter and this projection differ there), and
  * HAVAC hits that fall in no window.

usage: window_check.py <fasta> <hmm> <windows.tsv> [p-value] [device spec]
"""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from havac_b200 import havac  # noqa: E402


def read_windows(path):
    """readHmmerHits (hmmerSsvRef.cpp:379-413): header line skipped, malformed lines reported and skipped."""
    out = []
    with open(path) as fh:
        fh.readline()
        for line in fh:
            f = line.split()
            if len(f) < 6:
                if line.strip():
                    print(f"could not parse line: {line.rstrip()!r}", file=sys.stderr)
                continue
            out.append({"sequenceId": int(f[0], 0), "windowStart": int(f[1], 0), "windowEnd": int(f[2], 0),
                        "windowLength": int(f[3], 0), "hmmPosition": int(f[4], 0), "accession": f[5]})
    return out


def traceback_score(symbols, table, window_start, window_end, phmm_position):
    """attemptTraceback (hmmerSsvRef.cpp:504-586) on 0-based inclusive sequence coordinates inside
    `symbols` (one record's 2-bit symbols) and `table` (one model's int8 rows); `phmm_position` is the
    row under `window_end`. Returns the score at which the walk stopped (>= 256: verified)."""
    rows = len(table)
    length = window_end - window_start
    phmm_start = phmm_position - length
    if phmm_start < 0 or phmm_position >= rows or window_start < 0 or window_end >= len(symbols):
        return -1  # the window does not lie on a diagonal of this model
    acc = best = 0
    for step in range(length + 1):
        acc += int(table[phmm_start + step, symbols[window_start + step]])
        best = max(best, acc)
        if acc < 0:
            acc = 0
        if acc >= 256:
            return acc
    acc = best  # walk back from the window start
    for step in range(1, min(window_start, phmm_start) + 1):
        acc += int(table[phmm_start - step, symbols[window_start - step]])
        best = max(best, acc)
        if acc < 0:
            break
        if acc >= 256:
            return acc
    acc = best  # and forward from its end
    for step in range(1, min(len(symbols) - window_end, rows - phmm_position)):
        acc += int(table[phmm_position + step, symbols[window_end + step]])
        best = max(best, acc)
        if acc < 0:
            break
        if acc >= 256:
            return acc
    return best


def check(fasta, hmm, windows, p_value=0.02, spec="", seed=1):
    h = havac.Havac(0, p_value, spec)
    try:
        havac.srand(seed)
        h.loadPhmm(hmm)
        h.loadSequence(fasta)
        h.runHardwareClient()
        pos, seq_idx, phmm_pos, phmm_idx = h.getHitArraysFromFinishedRun()
        packed = h.devicePackedSequence()
        ends = h.sequenceEndPositions().astype(np.int64)
        table = h.devicePhmmScores()
    finally:
        h.close()
    _, lens = havac.project_phmm(hmm, p_value)
    names = havac.model_names(hmm)  # [(NAME, ACC)]
    first_row = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    starts = np.concatenate([[0], ends[:-1]])
    symbols = np.zeros(len(packed) * 4, dtype=np.uint8)
    for k in range(4):
        symbols[k::4] = (packed >> (2 * k)) & 3

    # HAVAC hits by (record, model, diagonal)
    hit_diag = {}
    for p, s, pp, pi in zip(pos.tolist(), seq_idx.tolist(), phmm_pos.tolist(), phmm_idx.tolist()):
        hit_diag.setdefault((s, pi, p - pp), []).append(pp)
    used = set()
    report = {"windows": len(windows), "verified": 0, "verified_and_reported": 0, "verified_but_missed": [], "below_threshold": 0,
              "off_model": 0, "lowest_score": None}
    for w in windows:
        model = next((i for i, na in enumerate(names) if w["accession"] in na), 0 if len(names) == 1 else -1)
        rec = w["sequenceId"]
        if model < 0 or rec >= len(ends):
            report["off_model"] += 1
            continue
        rec_sym = symbols[starts[rec]:ends[rec]]
        tab = table[first_row[model]:first_row[model + 1]]
        ws, we, hp = w["windowStart"] - 1, w["windowEnd"] - 1, w["hmmPosition"] + 1  # as :187
        score = traceback_score(rec_sym, tab, ws, we, hp)
        if score < 0:
            report["off_model"] += 1
            continue
        if score >= 256:
            report["verified"] += 1
            key = (rec, model, we - hp)
            if key in hit_diag:
                report["verified_and_reported"] += 1
                used.add(key)
            else:
                report["verified_but_missed"].append(w)
        else:
            report["below_threshold"] += 1
            report["lowest_score"] = score if report["lowest_score"] is None else min(report["lowest_score"], score)
    report["havac_hits"] = int(len(pos))
    report["havac_hit_diagonals"] = len(hit_diag)
    report["havac_hit_diagonals_in_no_window"] = len(set(hit_diag) - used)
    return report


def main():
    if len(sys.argv) < 4:
        print(__doc__)
        return 2
    p = float(sys.argv[4]) if len(sys.argv) > 4 else 0.02
    spec = sys.argv[5] if len(sys.argv) > 5 else ""
    report = check(sys.argv[1], sys.argv[2], read_windows(sys.argv[3]), p, spec)
    print(json.dumps(report))
    return 1 if report["verified_but_missed"] else 0


if __name__ == "__main__":
    sys.exit(main())
