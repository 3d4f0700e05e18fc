"""Generates the committed golden vectors of tests/golden/ by running the UNMODIFIED reference
objects (oracle/_ref/libhavac_ref.so, compiled from /root/reference by oracle/Makefile) on small
deterministic inputs. Run here (the reference tree is mounted only in the build container):

    python tests/golden/make_golden.py

Outputs, per case <name>:
    <name>.fa, <name>.hmm      the inputs (written by havac_b200/synth.py, fixed seeds)
    <name>.npz                 packed   : SequencePreprocessor buffer after srand(SEED)
                               seq_end  : FastaVector sequenceEndPosition per record
                               table    : PhmmPreprocessor int8 table at P_VALUE
                               model_len: LENG per model
                               hit_pos, hit_row : softSsvThreshold256 over the whole padded buffer
                               res_*    : the hits resolved with fastaVectorGetLocalSequencePositionFromGlobal
                                          + the prefix-sum model lookup (host/Havac.cpp:145-187)
Also scale_factors.npz: findThreshold256ScalingFactor on a grid of (mu, lambda, MAXL, LENG, p).
"""
import ctypes as C
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import oracle_lib as O  # noqa: E402
from havac_b200 import synth  # noqa: E402

SEED = 20240
CASES = {
    # name: (record lengths, model lengths, p-value, planted copies per model)
    "two_models": ([30_000, 18_001], [120, 77], 0.02, 6),
    "three_records": ([5_000, 12_288, 7_211], [200], 0.05, 5),
}


def resolve_with_reference(L, pos, row, seq_end, model_len):
    sums = np.concatenate([[0], np.cumsum(model_len)]).astype(np.uint64)
    out = []
    for p, r in zip(pos.tolist(), row.tolist()):
        si, sp = C.c_uint64(0), C.c_uint64(0)
        if not L.ref_local_position(seq_end, len(seq_end), p, C.byref(si), C.byref(sp)):
            continue
        mi = int(np.searchsorted(sums, r, side="right")) - 1
        out.append((sp.value, si.value, r - int(sums[mi]), mi))
    return np.array(out, dtype=np.uint64).reshape(-1, 4)


def main():
    L = O.ref()
    if L is None:
        raise SystemExit("oracle/_ref/libhavac_ref.so is missing: run `make -C oracle` where /root/reference is mounted")
    for idx, (name, (rec_lens, model_lens, p, copies)) in enumerate(CASES.items()):
        models = synth.random_models(SEED + idx, model_lens)
        sym = synth.random_symbols(SEED + idx, int(sum(rec_lens)))
        synth.plant_homologs(SEED + idx, sym, models, copies, substitution=0.12)
        edges = np.concatenate([[0], np.cumsum(rec_lens)])
        recs = [sym[edges[i]:edges[i + 1]] for i in range(len(rec_lens))]
        fa, hmm = os.path.join(HERE, name + ".fa"), os.path.join(HERE, name + ".hmm")
        synth.write_fasta(fa, recs)
        synth.write_hmm(hmm, models)
        packed, seq_end, count = O.ref_load_fasta(fa, seed=SEED)
        table, model_len = O.ref_load_phmm(hmm, p)
        pos, row = O.soft_ssv(synth.unpack_2bit(packed), table, lib=L)
        res = resolve_with_reference(L, pos, row, seq_end, model_len)
        np.savez_compressed(os.path.join(HERE, name + ".npz"), packed=packed, seq_end=seq_end, table=table,
                            model_len=model_len, hit_pos=pos, hit_row=row, res=res, p_value=np.float32(p),
                            seed=np.uint32(SEED), count=np.uint64(count))
        print(f"{name}: {count} symbols, {len(table) // 4} rows, {len(pos)} raw hits, {len(res)} resolved")
    g = np.random.Generator(np.random.PCG64(SEED))
    grid = []
    for _ in range(400):
        mu, lam = g.uniform(-12, -5), g.uniform(0.6, 0.8)
        leng = int(g.integers(20, 5000))
        maxl = int(leng * g.uniform(1.0, 3.0)) + 1
        p = float(g.choice([1e-12, 1e-9, 1e-5, 0.001, 0.02, 0.05, 0.1, 0.5]))
        s = L.ref_scale_factor(np.float32(mu), np.float32(lam), maxl, leng, np.float32(p))
        grid.append((np.float32(mu), np.float32(lam), maxl, leng, np.float32(p), np.float32(s)))
    a = np.array(grid, dtype=np.float64)
    np.savez_compressed(os.path.join(HERE, "scale_factors.npz"), mu=a[:, 0].astype(np.float32), lam=a[:, 1].astype(np.float32),
                        maxl=a[:, 2].astype(np.uint32), leng=a[:, 3].astype(np.uint32), p=a[:, 4].astype(np.float32),
                        scale=a[:, 5].astype(np.float32))
    print("scale_factors: 400 points")


if __name__ == "__main__":
    main()
