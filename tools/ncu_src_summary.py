"""Summarise an `ncu --page source --csv` export (SASS view) in blocks of instructions:
samples, executed count, top stall reasons, and the memory/sync opcodes in the block.
    python tools/ncu_src_summary.py src.csv [block]"""
import csv
import sys
from collections import Counter

rows = list(csv.reader(open(sys.argv[1])))
blk = int(sys.argv[2]) if len(sys.argv) > 2 else 100
hdr = rows[1]
col = {h: i for i, h in enumerate(hdr)}
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
data = rows[2:]
tot = sum(int(r[col["# Samples"]] or 0) for r in data)
print("total samples", tot, "instructions", len(data))
for b0 in range(0, len(data), blk):
    part = data[b0:b0 + blk]
    smp = sum(int(r[col["# Samples"]] or 0) for r in part)
    if smp < tot * 0.004:
        continue
    ex = max(int(r[col["Instructions Executed"]] or 0) for r in part)
    st = Counter()
    ops = Counter()
    for r in part:
        for s in stalls:
            v = r[col[s]]
            if v and v != "0":
                st[s[6:]] += int(v)
        op = r[col["Source"]].split()
        op = [o for o in op if not o.startswith("@")]
        if op:
            o = op[0].split(".")[0]
            if o in ("LDG", "STG", "LDS", "STS", "SHFL", "WARPSYNC", "BAR", "MUFU", "ATOMG", "BRA", "CALL", "SYNCS", "UBLKCP", "CCTL", "LDC", "VOTE", "BSSY", "BSYNC"):
                ops[o] += 1
    top = ", ".join(f"{k}:{v}" for k, v in st.most_common(4))
    print(f"[{b0:5d}] samples {smp:6d} ({100.0*smp/tot:4.1f}%) exec/inst {ex:8d}  {top}   | " + " ".join(f"{k}{v}" for k, v in ops.most_common()))
