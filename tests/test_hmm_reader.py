"""The HMMER3 reader (havac_b200/host/PhmmParser.cpp) on its own: number parsing against libc's
strtof (== the reference's sscanf "%f", P7HmmReader/src/p7HmmReader.c:694-1064) and the threaded
split at model ends against the single-threaded read of the same file."""
import ctypes as C
import os

import numpy as np
import pytest

from havac_b200 import havac, synth

_libc = C.CDLL(None)
_libc.strtof.restype = C.c_float
_libc.strtof.argtypes = [C.c_char_p, C.c_void_p]


def strtof(s):
    return np.float32(_libc.strtof(s.encode(), None))


def hmm_with_emissions(path, rows_of_tokens, name="m", tail_newline=True):
    """One model whose match lines carry the given token strings verbatim."""
    n = len(rows_of_tokens)
    with open(path, "w") as fh:
        fh.write(f"HMMER3/f [x]\nNAME  {name}\nLENG  {n}\nMAXL  {n + 50}\nALPH  DNA\n")
        fh.write("STATS LOCAL MSV      -8.5000  0.71000\n")
        fh.write("HMM          A        C        G        T   \n            m->m     m->i     m->d     i->m     i->i     d->m     d->d\n")
        fh.write("  COMPO   1.38629  1.38629  1.38629  1.38629\n          1.38629  1.38629  1.38629  1.38629\n")
        fh.write("          0.02000  4.60517  4.60517  0.69315  0.69315  0.00000        *\n")
        for k, toks in enumerate(rows_of_tokens):
            fh.write(f"{k + 1:7d}   " + "  ".join(toks) + "      - - - - -\n")
            fh.write("          1.38629  1.38629  1.38629  1.38629\n")
            fh.write("          0.02000  4.60517  4.60517  0.69315  0.69315  0.69315  0.69315\n")
        fh.write("//\n" if tail_newline else "//")


def test_numbers_parse_like_strtof(tmp_path):
    g = synth.rng(81, 1)
    toks = []
    # HMMER's own spelling, many of them
    toks += [f"{x:.5f}" for x in g.exponential(1.5, size=60_000)]
    # other plain decimals: few and many digits, no fraction, leading zeros, signs
    for digits in range(1, 16):
        for _ in range(300):
            m = int(g.integers(0, 10 ** digits))
            frac = int(g.integers(0, digits + 1))
            s = f"{m:0{digits}d}"
            s = s[:digits - frac] + "." + s[digits - frac:] if frac else s
            toks.append(("-" if g.random() < 0.2 else "+" if g.random() < 0.05 else "") + s)
    # spellings that must go to strtof: exponents, 16+ digits, inf/nan, hex, trailing junk
    toks += ["1e-3", "2.5E+2", "1.17549435e-38", "1e-45", "3.4028235e38", "1e39", "0x1.8p1", "inf", "nan", "1.5abc",
             "12345678901234567890.5", "0.000000000000000000001", "16777217", "16777216.99999999", ".5", "5.", "0", "-0", "000"]
    # float midpoints spelled exactly: 1 + 2^-24 and neighbours, 2^24 + 1
    toks += ["1.000000059604644775390625", "1.00000005960464", "1.00000005960465", "1.00000017881393", "16777217.0",
             "0.100000001490116", "0.10000000149011612", "8388608.5", "8388609.5", "4194304.25", "4194304.75"]
    while len(toks) % 4:
        toks.append("1.0")
    rows = [toks[i:i + 4] for i in range(0, len(toks), 4)]
    path = tmp_path / "numbers.hmm"
    hmm_with_emissions(path, rows)
    e = havac.parse_phmm(str(path)).reshape(-1)
    want = np.array([strtof(t) for t in toks], dtype=np.float32)
    same = (e.view(np.uint32) == want.view(np.uint32)) | (np.isnan(e) & np.isnan(want))
    assert same.all(), [(toks[i], e[i], want[i]) for i in np.flatnonzero(~same)[:10]]


def test_star_is_infinite_and_last_line_may_lack_a_newline(tmp_path):
    path = tmp_path / "star.hmm"
    hmm_with_emissions(path, [["*", "0.5", "1.25", "3"], ["1", "2", "*", "4"]], tail_newline=False)
    e = havac.parse_phmm(str(path))
    assert np.isposinf(e[0, 0]) and np.isposinf(e[1, 2]) and e[0, 1] == 0.5 and e[1, 3] == 4.0


def test_threaded_read_equals_single_threaded(tmp_path, monkeypatch):
    """~30 MB of models: ranges cut at "//" lines and parsed in parallel == one pass."""
    lens = synth.model_lengths(82, 420, 100, 700)
    models = synth.random_models(82, [int(x) for x in lens])
    path = str(tmp_path / "many.hmm")
    synth.write_hmm(path, models)
    assert os.path.getsize(path) > 24 << 20
    monkeypatch.setenv("HAVAC_PARSE_THREADS", "1")
    one = havac.parse_phmm(path)
    t1, l1 = havac.project_phmm(path, 0.02)
    for threads in ("2", "7", "16"):
        monkeypatch.setenv("HAVAC_PARSE_THREADS", threads)
        many = havac.parse_phmm(path)
        tn, ln = havac.project_phmm(path, 0.02)
        assert np.array_equal(one.view(np.uint32), many.view(np.uint32))
        assert np.array_equal(l1, ln) and np.array_equal(t1, tn)
    assert list(l1) == [int(x) for x in lens]


def test_threaded_read_keeps_the_errors(tmp_path, monkeypatch):
    """A broken model in the middle of a large file, a truncated last model and a '//' line inside
    a header give what the single pass gives."""
    models = synth.random_models(83, [600] * 260)
    good = str(tmp_path / "good.hmm")
    synth.write_hmm(good, models)
    text = open(good).read()
    assert len(text) > 12 << 20
    monkeypatch.setenv("HAVAC_PARSE_THREADS", "3")

    mid = text.index("\n    300 ", len(text) // 2)  # drop one match line of a model in the middle
    broken = tmp_path / "broken.hmm"
    broken.write_text(text[:mid + 1] + text[text.index("\n", mid + 1) + 1:])
    with pytest.raises(RuntimeError, match="not formatted correctly"):
        havac.parse_phmm(str(broken))

    cut = tmp_path / "cut.hmm"
    cut.write_text(text[:text.rindex("\n    590 ")])
    with pytest.raises(RuntimeError, match="not formatted correctly"):
        havac.parse_phmm(str(cut))

    # lines that begin with "//" where the single pass does not take them for a model end: an unknown
    # header tag, and insert-emission lines (only their token count is looked at). The cuts at 1/3
    # and 2/3 of the file land on such lines, the ranges do not end between models, and the file
    # is read again in one pass.
    at = text.index("\nLENG", len(text) // 2)
    odd_text = text[:at + 1] + "// not the end of a model\n" + text[at + 1:]
    insert_line = "          1.38629  1.38629  1.38629  1.38629\n"
    a, b = len(odd_text) // 4, 3 * len(odd_text) // 4
    odd_text = odd_text[:a] + odd_text[a:b].replace(insert_line, "//        1.38629  1.38629  1.38629  1.38629\n") + odd_text[b:]
    odd = tmp_path / "odd.hmm"
    odd.write_text(odd_text)
    many = havac.parse_phmm(str(odd))
    monkeypatch.setenv("HAVAC_PARSE_THREADS", "1")
    assert np.array_equal(many.view(np.uint32), havac.parse_phmm(str(odd)).view(np.uint32))
    assert np.array_equal(many.view(np.uint32), havac.parse_phmm(good).view(np.uint32))
