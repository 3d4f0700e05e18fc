"""Deterministic synthetic inputs for the SSV path.

The reference regenerates every fixture with ``srand(time(0))`` and an external
``hmmbuild`` binary (device/test/softwareTestbench.cpp:49-56,
test/generator/hmmSeqGenerator.cpp:72-121), neither of which is reproducible
offline. This module writes the same two file kinds -- a DNA FASTA and a
HMMER3 ``ALPH DNA`` profile file carrying exactly the fields
P7HmmReader/src/p7HmmReader.c parses -- from a fixed seed.

PRNG: numpy's PCG64 (bit-stream stable across numpy versions), master seed
0x4841564143 ("HAVAC"), one sub-stream per purpose.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import List, Optional, Sequence, Tuple

import numpy as np

MASTER_SEED = 0x4841564143
SEGMENT_SYMBOLS = 12288  # device/PublicDefines.h:18-22
_ALPHABET = np.frombuffer(b"ACGT", dtype=np.uint8)


def rng(*stream: int) -> np.random.Generator:
    return np.random.Generator(np.random.PCG64([MASTER_SEED, *stream]))


@dataclass
class Model:
    name: str
    length: int
    max_length: int
    mu: float
    lam: float
    probs: np.ndarray  # (length, 4) emission probabilities, rows sum to 1
    consensus: np.ndarray  # (length,) uint8 in 0..3

    def neg_log_emissions(self) -> np.ndarray:
        """Match emissions as the file stores them: -ln P, 5 decimals."""
        return np.round(-np.log(self.probs), 5)


def random_model(g: np.random.Generator, length: int, name: str) -> Model:
    consensus = g.integers(0, 4, size=length, dtype=np.uint8)
    pc = g.uniform(0.55, 0.95, size=length)
    probs = np.repeat(((1.0 - pc) / 3.0)[:, None], 4, axis=1)
    probs[np.arange(length), consensus] = pc
    mu = float(g.uniform(-9.5, -7.5))
    lam = float(g.uniform(0.69, 0.73))
    return Model(name, length, int(math.ceil(1.5 * length)), mu, lam, probs, consensus)


def random_models(seed: int, lengths: Sequence[int]) -> List[Model]:
    g = rng(seed, 1)
    return [random_model(g, int(L), f"synth{idx:05d}") for idx, L in enumerate(lengths)]


def model_lengths(seed: int, count: int, lo: int, hi: int) -> np.ndarray:
    return rng(seed, 2).integers(lo, hi + 1, size=count)


def _f(x: float) -> str:
    return f"{x:.5f}"


def write_hmm(path: str, models: Sequence[Model]) -> None:
    """HMMER3/f text, ALPH DNA, no optional annotations.

    Field requirements follow p7HmmReader.c: format tag :99, LENG :191-206,
    MAXL :207-222, ALPH :234-236, RF/MM/CONS/CS/MAP flags :253-357, STATS
    :495-534, HMM header + skipped label line :535-568, COMPO/insert0/begin
    transitions :572-682, per-node triplets :694-1064, terminator :687-692.
    """
    with open(path, "w") as fh:
        for m in models:
            nle = m.neg_log_emissions()
            fh.write("HMMER3/f [3.3.2 | Nov 2020]\n")
            fh.write(f"NAME  {m.name}\n")
            fh.write(f"LENG  {m.length}\n")
            fh.write(f"MAXL  {m.max_length}\n")
            fh.write("ALPH  DNA\n")
            fh.write("RF    no\nMM    no\nCONS  no\nCS    no\nMAP   no\n")
            fh.write("NSEQ  1\nEFFN  1.000000\nCKSUM 0\n")
            fh.write(f"STATS LOCAL MSV      {m.mu:.4f}  {m.lam:.5f}\n")
            fh.write(f"STATS LOCAL VITERBI  {m.mu - 0.5:.4f}  {m.lam:.5f}\n")
            fh.write(f"STATS LOCAL FORWARD  {-4.0:.4f}  {m.lam:.5f}\n")
            fh.write("HMM          A        C        G        T   \n")
            fh.write("            m->m     m->i     m->d     i->m     i->i     d->m     d->d\n")
            q = _f(-math.log(0.25))
            fh.write(f"  COMPO   {q}  {q}  {q}  {q}\n")
            fh.write(f"          {q}  {q}  {q}  {q}\n")
            fh.write("          0.02000  4.60517  4.60517  0.69315  0.69315  0.00000        *\n")
            for k in range(m.length):
                e = nle[k]
                fh.write(f"{k + 1:7d}   {_f(e[0])}  {_f(e[1])}  {_f(e[2])}  {_f(e[3])}      - - - - -\n")
                fh.write(f"          {q}  {q}  {q}  {q}\n")
                if k + 1 < m.length:
                    fh.write("          0.02000  4.60517  4.60517  0.69315  0.69315  0.69315  0.69315\n")
                else:
                    fh.write("          0.01000  4.60517        *  0.69315  0.69315  0.00000        *\n")
            fh.write("//\n")


def random_symbols(seed: int, n: int) -> np.ndarray:
    """i.i.d. uniform symbols 0..3, one per byte."""
    g = rng(seed, 3)
    out = np.empty(n, dtype=np.uint8)
    step = 1 << 26
    for a in range(0, n, step):
        b = min(n, a + step)
        raw = g.integers(0, 256, size=(b - a + 3) // 4, dtype=np.uint8)
        quad = np.stack([raw & 3, (raw >> 2) & 3, (raw >> 4) & 3, raw >> 6], axis=1).reshape(-1)
        out[a:b] = quad[: b - a]
    return out


def plant_homologs(seed: int, symbols: np.ndarray, models: Sequence[Model], copies: int,
                   substitution: float = 0.15) -> List[Tuple[int, int]]:
    """Overwrite ``copies`` windows per model with its consensus, ``substitution``
    of positions replaced by a uniform base. Returns (model index, start)."""
    g = rng(seed, 4)
    placed = []
    n = len(symbols)
    for mi, m in enumerate(models):
        if m.length >= n:
            continue
        for _ in range(copies):
            start = int(g.integers(0, n - m.length))
            seq = m.consensus.copy()
            mut = g.random(m.length) < substitution
            seq[mut] = g.integers(0, 4, size=int(mut.sum()), dtype=np.uint8)
            symbols[start:start + m.length] = seq
            placed.append((mi, start))
    return placed


def write_fasta(path: str, records: Sequence[np.ndarray], names: Optional[Sequence[str]] = None,
                width: int = 80) -> None:
    with open(path, "wb") as fh:
        for idx, sym in enumerate(records):
            name = names[idx] if names else f"seq{idx}"
            fh.write(b">" + name.encode() + b"\n")
            chars = _ALPHABET[sym]
            full = (len(chars) // width) * width
            if full:
                body = np.empty((full // width, width + 1), dtype=np.uint8)
                body[:, :width] = chars[:full].reshape(-1, width)
                body[:, width] = 10
                fh.write(body.tobytes())
            if full < len(chars):
                fh.write(chars[full:].tobytes() + b"\n")


def split_records(symbols: np.ndarray, n_records: int) -> List[np.ndarray]:
    """Cut into n_records pieces of nearly equal size (views, no copy)."""
    edges = np.linspace(0, len(symbols), n_records + 1).astype(np.int64)
    return [symbols[edges[i]:edges[i + 1]] for i in range(n_records)]


def pack_2bit(symbols: np.ndarray, pad_to_segment: bool = True) -> np.ndarray:
    """Symbol k -> byte k/4, bits (k%4)*2, LSB first; zero ('A') padding to a
    multiple of 12288 symbols (host/sequence/SequencePreprocessor.cpp:15-19,37-59)."""
    n = len(symbols)
    total = ((n + SEGMENT_SYMBOLS - 1) // SEGMENT_SYMBOLS) * SEGMENT_SYMBOLS if pad_to_segment else ((n + 3) // 4) * 4
    out = np.zeros(total // 4, dtype=np.uint8)
    full = (n // 4) * 4
    q = symbols[:full].reshape(-1, 4)
    out[: full // 4] = q[:, 0] | (q[:, 1] << 2) | (q[:, 2] << 4) | (q[:, 3] << 6)
    for k in range(full, n):
        out[k >> 2] |= np.uint8(int(symbols[k]) << ((k & 3) * 2))
    return out


def unpack_2bit(packed: np.ndarray, n_symbols: Optional[int] = None) -> np.ndarray:
    quad = np.stack([packed & 3, (packed >> 2) & 3, (packed >> 4) & 3, packed >> 6], axis=1).reshape(-1)
    return quad if n_symbols is None else quad[:n_symbols]


def random_packed(seed: int, n_symbols: int) -> np.ndarray:
    """Packed i.i.d. uniform DNA, padded to whole segments -- the bench-only
    injection path for sizes where a FASTA round trip would dominate."""
    total = ((n_symbols + SEGMENT_SYMBOLS - 1) // SEGMENT_SYMBOLS) * SEGMENT_SYMBOLS
    out = np.zeros(total // 4, dtype=np.uint8)
    g = rng(seed, 5)
    nb = n_symbols // 4
    step = 1 << 27
    for a in range(0, nb, step):
        b = min(nb, a + step)
        out[a:b] = g.integers(0, 256, size=b - a, dtype=np.uint8)
    for k in range(nb * 4, n_symbols):
        out[k >> 2] |= np.uint8(int(g.integers(0, 4)) << ((k & 3) * 2))
    return out


def consensus_score_table(seed: int, n_rows: int, hi: Tuple[int, int] = (14, 22),
                          lo: Tuple[int, int] = (-32, -9)) -> Tuple[np.ndarray, np.ndarray]:
    """An int8 [row][A,C,G,T] score table shaped like a projected nucleotide model (one
    positive consensus score per row, three negative ones; PhmmReprojection.cpp:109-145 at
    p = 0.02 gives these ranges for the random models above) without going through a .hmm
    file. Returns (table, consensus symbols)."""
    g = rng(seed, 77)
    t = g.integers(lo[0], lo[1] + 1, size=(n_rows, 4)).astype(np.int8)
    cons = g.integers(0, 4, size=n_rows)
    t[np.arange(n_rows), cons] = g.integers(hi[0], hi[1] + 1, size=n_rows)
    return t, cons.astype(np.uint8)


def plant_consensus(seed: int, symbols: np.ndarray, consensus: np.ndarray, copies: int,
                    substitution: float = 0.1, min_len: int = 40) -> int:
    """Overwrite ``copies`` windows of ``symbols`` with mutated fragments of ``consensus``
    (the role of test/generator/hmmSeqGenerator.cpp:72-84). Returns how many were placed."""
    g = rng(seed, 78)
    n, L = len(symbols), len(consensus)
    placed = 0
    for _ in range(copies):
        a = int(g.integers(0, L - min_len)) if L > 2 * min_len else 0
        b = int(g.integers(a + min_len, L + 1)) if L > 2 * min_len else L
        frag = consensus[a:b].copy()
        mut = g.random(len(frag)) < substitution
        frag[mut] = g.integers(0, 4, size=int(mut.sum()))
        if len(frag) >= n:
            continue
        at = int(g.integers(0, n - len(frag)))
        symbols[at:at + len(frag)] = frag
        placed += 1
    return placed
