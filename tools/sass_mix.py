#!/usr/bin/env python
"""Instruction mix of the innermost loop of a kernel in the built library (tuning aid)."""
import collections, re, subprocess, sys
lib, pat = sys.argv[1], sys.argv[2]
names = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
blocks = re.split(r"\n\s*Function : ", names)
for b in blocks[1:]:
    name = b.split("\n", 1)[0]
    if not re.search(pat, name):
        continue
    ins = [(int(m.group(1), 16), m.group(2)) for m in re.finditer(r"/\*([0-9a-f]{4})\*/\s+(.*?);", b)]
    back = [(a, int(re.search(r"0x([0-9a-f]+)", i).group(1), 16)) for a, i in ins if "BRA" in i and re.search(r"0x([0-9a-f]+)", i) and int(re.search(r"0x([0-9a-f]+)", i).group(1), 16) < a]
    if not back:
        continue
    a1, a0 = max(back, key=lambda x: x[0] - x[1])
    loop = [re.sub(r"^@!?U?P\d+\s+", "", i).split()[0].split(".")[0] for a, i in ins if a0 <= a <= a1]
    c = collections.Counter(loop)
    alu = sum(v for k, v in c.items() if k in ("MOV", "FSEL", "FSETP", "FMNMX", "FMNMX3", "IADD3", "LOP3", "SHF", "ISETP", "SEL", "PLOP3", "IMNMX", "LEA", "P2R", "R2P", "DSETP", "IADD", "VIADD", "PRMT"))
    fma = sum(v for k, v in c.items() if k in ("FFMA", "FMUL", "FADD", "IMAD", "HFMA2", "FFMA2", "FMUL2", "FADD2"))
    print(name[:70], "total", len(loop), "alu~", alu, "fma~", fma)
    print("   ", dict(c.most_common(16)))
