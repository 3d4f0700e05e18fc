"""loadSequence wall time vs the number of staging threads (HAVAC_INGEST_THREADS) on a synthetic
FASTA; appends JSON lines to gpurun_out/ingest_sweep.jsonl.  usage: ingest_sweep.py [Mbp]"""
import json
import os
import sys
import tempfile
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from havac_b200 import havac, synth  # noqa: E402

mbp = int(sys.argv[1]) if len(sys.argv) > 1 else 100
with tempfile.TemporaryDirectory() as tmp:
    fa = os.path.join(tmp, "s.fa")
    sym = synth.random_symbols(5, mbp * 1_000_000)
    synth.write_fasta(fa, synth.split_records(sym, 4))
    del sym
    size = os.path.getsize(fa)
    out = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gpurun_out")
    os.makedirs(out, exist_ok=True)
    for threads in ("0", "1", "2", "4", "8"):  # 0: the library's own choice
        if threads == "0":
            os.environ.pop("HAVAC_INGEST_THREADS", None)
        else:
            os.environ["HAVAC_INGEST_THREADS"] = threads
        times = []
        for rep in range(4):
            h = havac.Havac(0, 0.02)
            havac.srand(1)
            t0 = time.perf_counter()
            h.loadSequence(fa)
            times.append(time.perf_counter() - t0)
            h.close()
        line = {"mbp": mbp, "file_bytes": size, "threads": int(threads), "best_s": min(times), "all_s": times,
                "GBps": size / min(times) / 1e9}
        print(json.dumps(line))
        with open(os.path.join(out, "ingest_sweep.jsonl"), "a") as fh:
            fh.write(json.dumps(line) + "\n")
