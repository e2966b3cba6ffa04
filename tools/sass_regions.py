"""Static view of a kernel's SASS (cuobjdump -sass -fun <mangled> lib.so > x.sass): regions split at RET / EXIT
(= the __noinline__ functions), instruction counts and marker mnemonics.  usage: sass_regions.py x.sass [region]"""
import re
import sys
from collections import Counter

MARKS = ("UBLKCP", "SYNCS", "LDGSTS", "MUFU", "SHFL", "LDL", "STL", "DFMA", "DMUL", "DADD", "MEMBAR", "LDS", "STS",
         "F2I", "FRND", "BRA", "UMOV", "LDC")
ins = []
for l in open(sys.argv[1]):
    m = re.search(r"/\*([0-9a-f]+)\*/\s+(.*?);", l)
    if m:
        ins.append((int(m.group(1), 16), m.group(2)))
reg, cur = [], []
for a, s in ins:
    cur.append((a, s))
    if re.search(r"\bRET\b|\bEXIT\b", s):
        reg.append(cur)
        cur = []
if cur:
    reg.append(cur)
print("static instructions", len(ins))
if len(sys.argv) > 2:
    for a, s in reg[int(sys.argv[2])]:
        print(hex(a), s)
else:
    for i, r in enumerate(reg):
        c = Counter()
        for a, s in r:
            for m in MARKS:
                if re.search(r"\b" + m, s):
                    c[m] += 1
        print(i, hex(r[0][0]), len(r), dict(c))
