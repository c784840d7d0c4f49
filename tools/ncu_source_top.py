"""Summarise an `ncu --page source --csv` export of the round kernel: samples per code region (spin
waits apart), and the hottest SASS lines with their main stall reasons."""
import csv
import sys

R = list(csv.reader(open(sys.argv[1])))
H, rows = R[1], R[2:]
ix = {n: i for i, n in enumerate(H)}
stalls = [k for k in H if k.startswith("stall_") and "Not Issued" not in k]
tot = sum(int(r[4]) for r in rows)


def top(r, n=2):
    st = {k: int(r[ix[k]] or 0) for k in stalls}
    ks = sorted(st, key=st.get, reverse=True)[:n]
    return " ".join(f"{k[6:]}:{st[k]}" for k in ks if st[k])


print(R[0][1][:120], "total samples", tot)
spin = [i for i, r in enumerate(rows) if "BRA" in r[1] and int(r[4]) > 0.002 * tot and top(r, 1).startswith("long_sb")]
sp = sum(int(rows[i][4]) for i in spin)
print(f"spin-wait branches: {100 * sp / tot:.1f}% of samples at rows {spin}")
nshow = int(sys.argv[2]) if len(sys.argv) > 2 else 40
order = sorted(range(len(rows)), key=lambda i: -int(rows[i][4]))[:nshow]
for i in sorted(order):
    r = rows[i]
    print(f"{i:5d} {r[1][:60]:60s} {int(r[4]):7d} {100 * int(r[4]) / tot:5.1f}% {r[5]:>10s} {top(r)}")
print("samples per 100 SASS rows:")
for a in range(0, len(rows), 100):
    s = sum(int(r[4]) for r in rows[a:a + 100])
    nd = sum(1 for r in rows[a:a + 100] if "DMMA" in r[1])
    if s > 0.004 * tot:
        print(f"  {a:5d}: {100 * s / tot:5.1f}%  ({nd} DMMA)")
