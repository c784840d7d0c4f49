"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per kernel the launch count,
total and mean duration, and its share of all kernel time; optionally only the first `--fits N`
worth of launches of a run whose first fit has `--per-fit L` launches."""
import collections
import csv
import re
import sys

path = sys.argv[1]
first = int(sys.argv[2]) if len(sys.argv) > 2 else None   # only the first N launches (e.g. one fit)
rows = [r for r in csv.reader(open(path)) if len(r) > 10]
H = rows[0]
ix = {n: i for i, n in enumerate(H)}
tot = collections.OrderedDict()
n = 0
for r in rows[1:]:
    if r[ix["Metric Name"]] != "gpu__time_duration.sum":
        continue
    n += 1
    if first is not None and n > first:
        break
    name = re.sub(r"\(.*", "", r[ix["Kernel Name"]])
    name = re.sub(r"<unnamed>::|\(anonymous namespace\)::|void ", "", name)
    v = float(r[ix["Metric Value"]].replace(",", ""))
    unit = r[ix["Metric Unit"]]
    ms = v / 1e6 if unit in ("ns", "nsecond") else v / 1e3 if unit in ("us", "usecond") else v if unit in ("ms", "msecond") else v * 1e3
    c = tot.setdefault(name, [0, 0.0])
    c[0] += 1
    c[1] += ms
total = sum(c[1] for c in tot.values())
print(f"# {path}: {n if first is None else min(n, first)} launches, {total:.1f} ms of kernel time (ncu: serialised, cold caches)")
print(f"{'kernel':70s} {'launches':>8s} {'total ms':>10s} {'mean us':>10s} {'share':>7s}")
for name, (cnt, ms) in sorted(tot.items(), key=lambda kv: -kv[1][1]):
    print(f"{name[:70]:70s} {cnt:8d} {ms:10.2f} {1e3 * ms / cnt:10.1f} {100 * ms / total:6.1f}%")
