"""Top stall sites of an ncu source page (sass view).  python tools/ncu_hot.py src.csv [N]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 40
hdr = rows[1]
ia, isrc, iall = hdr.index("Address"), hdr.index("Source"), hdr.index("Warp Stall Sampling (All Samples)")
stall_cols = [(i, h) for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
body = []
for r in rows[2:]:
    if len(r) != len(hdr) or r[ia] == "Address":
        break
    body.append(r)
tot = sum(int(r[iall] or 0) for r in body)
print("total samples", tot)
order = sorted(range(len(body)), key=lambda i: -int(body[i][iall] or 0))[:n]
for i in sorted(order):
    r = body[i]
    st = sorted(((int(r[c] or 0), h) for c, h in stall_cols), reverse=True)[:2]
    print(f"{i:5d} {int(r[iall]):6d} {100*int(r[iall])/tot:5.1f}%  {r[isrc].strip()[:90]:90s} {st}")
