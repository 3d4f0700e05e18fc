"""Development aid: opcode pattern of the fast kernel's hot loop in a built library.
Prints how many PRMT-PRMT / HFMA2-HFMA2 neighbour pairs the loop has."""
import re, subprocess, sys
lib = sys.argv[1]
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
m = re.search(r"Function : \S*ssv_f16x2_kernel.*?(?=Function :|\Z)", sass, re.S)
ops = re.findall(r"/\*[0-9a-f]{4,5}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", m.group(0))
s = "".join("P" if o.startswith("PRMT") else "H" if o.startswith("HFMA2") else "L" if o.startswith("LDS") else
            "M" if "MOV" in o else "B" if o.startswith("BRA") else "." for o in ops)
best = max(re.finditer(r"[PHLM.]{400,}", s), key=lambda x: x.group(0).count("P"), default=None)
seg = best.group(0) if best else s
first, last = seg.find("P"), seg.rfind("H")
seg = seg[first:last + 1]
print(f"{lib}: loop {len(seg)} instr  P={seg.count('P')} H={seg.count('H')} L={seg.count('L')} M={seg.count('M')} other={seg.count('.')}"
      f"  PP={len(re.findall('(?=PP)', seg))} HH={len(re.findall('(?=HH)', seg))}")
if "-v" in sys.argv:
    for k in range(0, len(seg), 100):
        print(seg[k:k + 100])
