"""Quick device-timed throughput probe (development aid; bench.py is the contract)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from havac_b200 import synth
from havac_b200 import build as _b
if os.environ.get('HAVAC_CUDA_LIB'):
    _b.CUDA_LIB = os.path.abspath(os.environ['HAVAC_CUDA_LIB'])
from havac_b200.client import HavacHwClient

def main():
    n = int(float(sys.argv[1])) if len(sys.argv) > 1 else 100_000_000
    rows = int(float(sys.argv[2])) if len(sys.argv) > 2 else 35_000
    reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
    pos_lo = int(sys.argv[4]) if len(sys.argv) > 4 else 14  # consensus scores pos_lo..pos_lo+8: raise for denser hits
    g = synth.rng(99, 1)
    t = g.integers(-32, -8, size=(rows, 4)).astype(np.int8)
    t[np.arange(rows), g.integers(0, 4, size=rows)] = g.integers(pos_lo, pos_lo + 9, size=rows)
    packed = synth.random_packed(99, n)
    c = HavacHwClient(";hitcap=200000000+")
    if os.environ.get("HAVAC_KERNEL"):
        c.setKernel(os.environ["HAVAC_KERNEL"])
    c.writeSequence(packed); c.writePhmm(t)
    for i in range(reps):
        t0 = time.time(); c.invokeHavacSsvAsync(); st = c.waitForHavacSsvAsync(); wall = time.time() - t0
        ms = c.deviceTimeMs()
        print(f"run {i}: state={st} hits={c.getNumHits()} device={ms:.2f} ms wall={wall*1e3:.1f} ms  GCUPS={n*rows/ms/1e6:.1f} effMHz={c.smCycles()/ms/1e3:.0f}", flush=True)

if __name__ == "__main__":
    main()
