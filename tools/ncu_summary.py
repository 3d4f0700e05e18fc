"""Turns `ncu -i X.ncu-rep --page raw --csv` (stdin) into the `name [unit] = value` lines kept
under profiles/ (development aid). Usage: ncu -i rep --page raw --csv | python tools/ncu_summary.py"""
import csv, sys

rows = [r for r in csv.reader(sys.stdin) if r]
hdr = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
names, units = rows[hdr], rows[hdr + 1]
for vals in rows[hdr + 2:]:
    print(f"# kernel: {vals[names.index('Kernel Name')]}  grid {vals[names.index('Grid Size')]} block {vals[names.index('Block Size')]}")
    out = []
    for n, u, v in zip(names, units, vals):
        if "." not in n or n in ("Process Name",):
            continue
        # drop ncu's section prefix (e.g. FBSP.TriageCompute.) where there is one
        parts = n.split(".")
        while parts and (parts[0].isupper() or parts[0] in ("TriageCompute", "TriageSCG", "TriageAC")) and "__" not in parts[0]:
            parts = parts[1:]
        out.append((".".join(parts), u, v))
    for n, u, v in sorted(set(out)):
        print(f"{n} [{u}] = {v}" if u else f"{n} = {v}")
