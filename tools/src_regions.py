"""Where the warps of a kernel spend their time, by straight-line SEGMENT of its SASS (development aid).

Input: `ncu -i X.ncu-rep --page source --csv --print-source sass` (file argument or stdin). The listing is cut at every
control-flow instruction (branch, call, return, barrier, warp sync, exit); each segment is one line: address of its
first instruction, number of instructions, how often its first instruction ran, its stall samples (share of the
kernel's), the three commonest stall reasons, and the instruction it ends with. Segments below --min percent of the
samples are folded into one "others" line.

Usage: python tools/src_regions.py gpurun_out/r2_src_f16x2.csv [--min 0.5]"""
import csv
import sys

CUT = ("BRA", "BRX", "JMP", "CALL", "RET", "EXIT", "BAR.", "WARPSYNC", "BSYNC", "BSSY", "BREAK")


def main():
    args = [a for a in sys.argv[1:] if not a.startswith("--")]
    floor = float(sys.argv[sys.argv.index("--min") + 1]) if "--min" in sys.argv else 0.5
    rows = list(csv.reader(open(args[0]) if args else sys.stdin))
    hdr = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    names = rows[hdr]
    col = {n: i for i, n in enumerate(names)}
    stall_cols = [(n, i) for i, n in enumerate(names) if n.startswith("stall_") and "Not Issued" not in n]
    segs, cur = [], None
    for r in rows[hdr + 1:]:
        if len(r) < len(names) or not r[0].startswith("0x"):
            continue
        text = r[col["Source"]].strip()
        samples = int(r[col["# Samples"]] or 0)
        if cur is None:
            cur = {"addr": r[0], "n": 0, "exec": int(r[col["Instructions Executed"]] or 0), "samples": 0, "stalls": {}, "last": ""}
        cur["n"] += 1
        cur["samples"] += samples
        cur["last"] = text
        for n, i in stall_cols:
            v = int(r[i] or 0)
            if v:
                cur["stalls"][n] = cur["stalls"].get(n, 0) + v
        op = text.split()[1] if text.startswith("@") and len(text.split()) > 1 else (text.split()[0] if text else "")
        if op.startswith(CUT):
            segs.append(cur)
            cur = None
    if cur:
        segs.append(cur)
    total = sum(s["samples"] for s in segs) or 1
    print(f"# {len(segs)} segments, {total} samples; segments under {floor} % folded")
    print(f"# {'first address':>16} {'instr':>6} {'runs of 1st instr':>18} {'samples':>9} {'share':>7}  top stalls | ends with")
    rest = 0
    for s in segs:
        share = 100.0 * s["samples"] / total
        if share < floor:
            rest += s["samples"]
            continue
        top = sorted(s["stalls"].items(), key=lambda kv: -kv[1])[:3]
        tops = " ".join(f"{k[6:]}={100.0 * v / max(1, sum(s['stalls'].values())):.0f}%" for k, v in top)
        print(f"  {s['addr']:>16} {s['n']:>6} {s['exec']:>18} {s['samples']:>9} {share:6.2f}%  {tops} | {s['last'][:60]}")
    print(f"  {'others':>16} {'':>6} {'':>18} {rest:>9} {100.0 * rest / total:6.2f}%")


if __name__ == "__main__":
    main()
