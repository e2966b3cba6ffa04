"""Per-function sample table from an ncu source page.

usage: ncu_regions.py <source_sass.csv> [title]
Input: `ncu -i X.ncu-rep --page source --csv --print-source sass` of one kernel launch captured with
`--set full --import-source on`.  The SASS listing is split at RET / EXIT (= the __noinline__ functions the
kernel is made of); for every region: static instruction count, share of executed warp instructions, share of
stall samples, the dominant stall reasons and a few marker mnemonics to recognise the function by.
"""
import csv
import sys
from collections import Counter

MARKS = ("UBLKCP", "SYNCS", "LDGSTS", "MUFU", "SHFL", "LDL", "STL", "DFMA", "MEMBAR", "ATOMG", "LDS", "STS")


def main(path, title=""):
    rows = []
    with open(path, newline="") as fh:
        header = None
        for r in csv.reader(fh):
            if header is None:
                if r and r[0] == "Address":
                    header = r
                continue
            if len(r) >= len(header) - 2:
                rows.append(r)
    col = {h: i for i, h in enumerate(header)}
    stall_cols = [h for h in header if h.startswith("stall_") and "Not Issued" not in h]

    def num(r, h):
        try:
            return float(r[col[h]])
        except (ValueError, IndexError):
            return 0.0

    regions, cur = [], []
    for r in rows:
        cur.append(r)
        op = r[col["Source"]]
        if " RET" in op or op.strip().startswith("RET") or "EXIT" in op:
            regions.append(cur)
            cur = []
    if cur:
        regions.append(cur)
    tot_s = sum(num(r, "# Samples") for r in rows) or 1.0
    tot_i = sum(num(r, "Instructions Executed") for r in rows) or 1.0
    print(title)
    print(f"total samples {int(tot_s)}, warp instructions executed {int(tot_i)}, static instructions {len(rows)}")
    out = []
    for idx, reg in enumerate(regions):
        s = sum(num(r, "# Samples") for r in reg)
        i = sum(num(r, "Instructions Executed") for r in reg)
        st = Counter()
        for h in stall_cols:
            st[h.replace("stall_", "")] = sum(num(r, h) for r in reg)
        stot = sum(st.values()) or 1.0
        top = [(k, round(v / stot, 2)) for k, v in st.most_common(4)]
        marks = [m for m in MARKS if any(m in r[col["Source"]] for r in reg)]
        out.append((s, idx, len(reg), i, top, marks))
    for s, idx, n, i, top, marks in sorted(out, reverse=True)[:24]:
        print(f"region {idx:3d} static {n:5d} executed {100 * i / tot_i:5.1f}% samples {100 * s / tot_s:5.1f}%  "
              f"instr/sample {i / max(s, 1):8.0f}  top stalls {top} contains {marks}")


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else "")
