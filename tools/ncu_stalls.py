"""Developer tool: summarise an `ncu --page source --csv` dump: top stalled SASS lines and stall reasons per address range.
Usage: python tools/ncu_stalls.py src.csv [lo:hi:name ...]   (addresses are hex offsets from the kernel start)"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
h, data = rows[1], rows[2:]
ia, isrc, isamp = h.index("Address"), h.index("Source"), h.index("# Samples")
iex = h.index("Instructions Executed")
stall_cols = [i for i, k in enumerate(h) if k.startswith("stall_") and "Not Issued" not in k]
addr = lambda r: int(r[ia], 16) if r[ia].startswith("0x") else int(r[ia])
base = addr(data[0])
tot = sum(int(r[isamp]) for r in data)
print("total samples", tot)
for spec in sys.argv[2:]:
    lo, hi, name = spec.split(":")
    lo, hi = int(lo, 16), int(hi, 16)
    s, st, n, ex = 0, {}, 0, 0
    for r in data:
        if lo <= addr(r) - base < hi:
            s += int(r[isamp]); n += 1; ex = max(ex, int(r[iex] or 0))
            for i in stall_cols:
                st[h[i]] = st.get(h[i], 0) + int(r[i] or 0)
    print(f"{name:12s} {s:7d} samples ({100 * s / tot:5.1f}%) {n:4d} instrs  ", sorted(st.items(), key=lambda kv: -kv[1])[:5])
for r in sorted(data, key=lambda r: -int(r[isamp]))[:int(28)]:
    st = sorted([(h[i], int(r[i] or 0)) for i in stall_cols], key=lambda kv: -kv[1])[:3]
    print(hex(addr(r) - base), r[isamp], r[iex], r[isrc][:70], st)
