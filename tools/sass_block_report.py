#!/usr/bin/env python
"""tools/sass_block_report.py LIB [KERNEL] -- the fast kernel's 32-row block as ptxas scheduled it.

Finds the longest run of PRMT / HFMA2.RELU / LDS instructions of `ssv_f16x2_kernel` in a built library
(cuobjdump -sass), prints every instruction of it with the scheduling-control fields decoded from the second
64-bit word (stall count, yield, write / read scoreboard, wait mask, operand-reuse flags), and the statistics
DESIGN.md section 6 argues from:

  * opcode mix and same-pipe neighbour pairs (PRMT-PRMT, HFMA2-HFMA2),
  * .reuse coverage of the PRMTs' two table registers,
  * register-bank histograms (index mod 2 and mod 4) of the three per-pair operands the register file must
    deliver when the table registers come from the reuse cache: the PRMT selector, the HFMA2 lookup operand
    and the HFMA2 state operand; and how often two or three of them share a bank,
  * issue distance (sum of stall counts) between each PRMT and the HFMA2 that consumes it.
"""
import collections
import re
import subprocess
import sys


def parse(lib, kernel):
    sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
    m = re.search(r"Function : \S*%s\S*(.*?)(?=Function :|\Z)" % re.escape(kernel), sass, re.S)
    if not m:
        raise SystemExit(f"{kernel} not found in {lib}")
    lines = m.group(1).split("\n")
    ins, i = [], 0
    while i < len(lines) - 1:
        a = re.match(r"\s+/\*([0-9a-f]{4,6})\*/\s+(.*?);\s+/\* (0x[0-9a-f]{16}) \*/", lines[i])
        b = re.match(r"\s+/\* (0x[0-9a-f]{16}) \*/", lines[i + 1])
        if a and b:
            ins.append({"addr": int(a.group(1), 16), "text": a.group(2).strip(), "w0": int(a.group(3), 16), "w1": int(b.group(1), 16)})
            i += 2
        else:
            i += 1
    for x in ins:
        w1 = x["w1"]
        x.update(stall=(w1 >> 41) & 0xF, yld=(w1 >> 45) & 1, wbar=(w1 >> 46) & 7, rbar=(w1 >> 49) & 7, wait=(w1 >> 52) & 0x3F,
                 reuse=(w1 >> 58) & 0xF)
        t = re.sub(r"^@!?U?P\d+\s+", "", x["text"])
        x["op"] = t.split()[0]
        x["regs"] = [int(r) for r in re.findall(r"\bR(\d+)\b", t)]
    return ins


def hot_block(ins):
    cls = "".join("P" if x["op"].startswith("PRMT") else "H" if x["op"].startswith("HFMA2") else
                  "L" if x["op"].startswith("LDS") else
                  "." if x["op"].startswith(("BRA", "BSSY", "BSYNC", "EXIT", "BAR", "STS", "STG", "ATOM", "RED", "CALL", "RET", "WARPSYNC",
                                             "VOTE", "SHFL", "LDG", "LDL", "STL", "SYNCS", "UBLKCP", "LDGSTS", "VIMNMX", "NOP", "DEPBAR",
                                             "S2R", "S2UR", "R2UR", "LDC", "ISETP", "ULEA", "UIADD3", "UMOV", "USEL", "UISETP", "UIMAD", "ULOP3",
                                             "USHF", "LEA", "SHF", "SEL", "IADD3")) else "o"
                  for x in ins)
    best = max(re.finditer(r"[PHLo]{300,}", cls), key=lambda m: m.group(0).count("P"), default=None)
    if best is None:
        raise SystemExit("no hot block found")
    a, b = best.start(), best.end()
    while not ins[a]["op"].startswith(("PRMT", "LDS")):
        a += 1
    while not ins[b - 1]["op"].startswith("HFMA2"):
        b -= 1
    return a, b


def main():
    lib = sys.argv[1]
    kernel = sys.argv[2] if len(sys.argv) > 2 else "ssv_f16x2_kernel"
    ins = parse(lib, kernel)
    a, b = hot_block(ins)
    blk = ins[a:b]
    ops = collections.Counter(x["op"] for x in blk)
    print(f"# {kernel} in {lib}: hot block {blk[0]['addr']:#x}..{blk[-1]['addr']:#x}, {len(blk)} instructions")
    print("# opcode mix:", dict(ops.most_common()))
    seq = "".join("P" if x["op"].startswith("PRMT") else "H" if x["op"].startswith("HFMA2") else "." for x in blk)
    print(f"# neighbour pairs: PRMT-PRMT {len(re.findall('(?=PP)', seq))}, HFMA2-HFMA2 {len(re.findall('(?=HH)', seq))}, "
          f"PRMT-HFMA2 or HFMA2-PRMT {len(re.findall('(?=PH|HP)', seq))}")
    runs = collections.Counter(len(m.group(0)) for m in re.finditer(r"P+|H+", seq))
    print("# run lengths of same-pipe instructions (length: count):", dict(sorted(runs.items())))
    prmt = [x for x in blk if x["op"].startswith("PRMT") and len(x["regs"]) == 4]  # (not the "PRMT R, RZ, imm, RZ" that makes a zero)
    hfma = [x for x in blk if x["op"].startswith("HFMA2")]
    both = sum(1 for x in prmt if x["reuse"] & 5 == 5)
    print(f"# PRMT .reuse on both table registers: {both} of {len(prmt)}; on none: {sum(1 for x in prmt if x['reuse'] & 5 == 0)}")
    # is the table pair of a PRMT the one the previous PRMT read (a reuse-cache hit needs the flag on THAT one)?
    hit = 0
    prev = None
    for x in prmt:
        pair = (x["regs"][1], x["regs"][3])
        if prev is not None and prev["reuse"] & 5 == 5 and (prev["regs"][1], prev["regs"][3]) == pair:
            hit += 1
        prev = x
    print(f"# PRMTs whose table pair was flagged .reuse by the PRMT before them (reuse-cache hits if the cache survives the "
          f"HFMA2 in between): {hit} of {len(prmt)}")
    for mod in (2, 4):
        sel = collections.Counter(x["regs"][2] % mod for x in prmt)
        mm = collections.Counter(x["regs"][1] % mod for x in hfma)
        ww = collections.Counter(x["regs"][2] % mod for x in hfma)
        print(f"# register index mod {mod}: PRMT selector {dict(sorted(sel.items()))}  HFMA2 first (slot A) operand {dict(sorted(mm.items()))}  "
              f"HFMA2 last (slot C) operand {dict(sorted(ww.items()))}")
        same = sum(1 for x in hfma if x["regs"][1] % mod == x["regs"][2] % mod)
        print(f"#   HFMA2 with both register operands in one bank (mod {mod}): {same} of {len(hfma)}")
        tri = sum(1 for x in prmt if len({r % mod for r in x["regs"][1:4]}) == 1)
        print(f"#   PRMT with all three source registers in one bank (mod {mod}): {tri} of {len(prmt)}")
    # producer -> consumer issue distance
    t, at, last_def = 0, {}, {}
    dist = collections.Counter()
    for x in blk:
        at[x["addr"]] = t
        if x["op"].startswith("HFMA2"):
            for src in x["regs"][1:]:  # the lookup result is the first or the last operand (fma(m, 1, w) == fma(w, 1, m))
                if src in last_def and last_def[src][1].startswith("PRMT"):
                    dist[t - last_def[src][0]] += 1
        if x["regs"] and not x["op"].startswith(("STS", "ST.")):
            last_def[x["regs"][0]] = (t, x["op"])
        t += max(1, x["stall"])
    print("# issue distance PRMT -> consuming HFMA2 in cycles by the stall counts (distance: count):", dict(sorted(dist.items())))
    print(f"# block length by stall counts: {t} cycles for {len(prmt) + len(hfma)} pair instructions + {len(blk) - len(prmt) - len(hfma)} others "
          f"=> {len(blk) / t:.3f} instructions per cycle if a warp ran alone and never waited")
    print("#")
    print("# addr  stall yield wbar rbar wait reuse  instruction")
    for x in blk:
        print(f"{x['addr']:05x}  s{x['stall']:<2d} y{x['yld']} w{x['wbar'] if x['wbar'] != 7 else '-'} r{x['rbar'] if x['rbar'] != 7 else '-'} "
              f"wt{x['wait']:02x} ru{x['reuse']:x}  {x['text']}")


if __name__ == "__main__":
    main()
