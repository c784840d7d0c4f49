"""Summarise ATE_PROFILE_CSV output (one line per timed launch) by Gram width."""
import collections, csv, sys
rows = list(csv.DictReader(open(sys.argv[1])))
p = int(sys.argv[2]) if len(sys.argv) > 2 else 256
b = collections.defaultdict(lambda: [0, 0.0, 0.0])
for r in rows:
    g = int(r["gram_cols"]); ms = float(r["ms"]); n = int(r["rows"])
    key = "plain-gram" if g < 0 else (g + 31) // 32 * 32
    b[key][0] += 1; b[key][1] += ms; b[key][2] += n
tot = sum(v[1] for v in b.values())
for k in sorted(b, key=lambda x: (isinstance(x, str), x)):
    v = b[k]
    print(f"cols<={k}: n={v[0]:5d} total={v[1]:8.1f} ms ({v[1]/tot*100:4.1f}%) avg={v[1]/v[0]:.3f} ms "
          f"avg rows={v[2]/v[0]:.0f}  GB/s={v[2]/v[0]*(p+2)*8/(v[1]/v[0]*1e-3)/1e9:.0f}")
print("launches", len(rows), "total ms", round(tot, 1))
