"""From an `ncu --set full --page raw --csv` export of irls_round_kernel launches: the per-launch figures the
bench line's roofline object quotes (DRAM bytes, pipes, shared-memory conflicts), and the JSON bench.py reads
(profiles/r02_irls_round_traffic.json)."""
import csv
import json
import sys

R = list(csv.reader(open(sys.argv[1])))
H, U = R[0], R[1]
ix = {n: i for i, n in enumerate(H)}
conv = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1, "Tbyte": 1e12}
tconv = {"ms": 1.0, "us": 1e-3, "ns": 1e-6, "s": 1e3, "msecond": 1.0, "usecond": 1e-3, "nsecond": 1e-6, "second": 1e3}


def g(row, name):
    v = row[ix[name]].replace(",", "")
    return float(v) if v not in ("", "n/a") else None


rows = []
for row in R[2:]:
    if "irls_round_kernel" not in row[ix["Kernel Name"]]:
        continue
    rd = g(row, "dram__bytes_read.sum") * conv[U[ix["dram__bytes_read.sum"]]]
    wr = g(row, "dram__bytes_write.sum") * conv[U[ix["dram__bytes_write.sum"]]]
    ms = g(row, "gpu__time_duration.sum") * tconv[U[ix["gpu__time_duration.sum"]]]
    wf = g(row, "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum")
    cf = g(row, "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum")
    rows.append(dict(kernel=row[ix["Kernel Name"]].replace("<unnamed>::", "")[:48], grid=row[ix["Grid Size"]], ms=ms,
                     dram_read_bytes=rd, dram_write_bytes=wr, dram_gbs=(rd + wr) / (ms * 1e-3) / 1e9,
                     dmma_pipe_pct=g(row, "sm__pipe_tensor_subpipe_dmma_cycles_active.avg.pct_of_peak_sustained_active"),
                     fp64_pipe_pct=g(row, "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active"),
                     issue_active_pct=g(row, "smsp__issue_active.avg.pct_of_peak_sustained_active"),
                     l2_hit_pct=g(row, "lts__t_sector_hit_rate.pct"),
                     shared_wavefronts=wf, shared_bank_conflicts=cf, shared_conflict_share=cf / wf if wf else None))
n = len(rows)
avg = {k: sum(r[k] for r in rows) / n for k in rows[0] if isinstance(rows[0][k], float)}
out = dict(source=sys.argv[1], what=sys.argv[2] if len(sys.argv) > 2 else "", launches_captured=n,
           dram_bytes_per_launch=avg["dram_read_bytes"] + avg["dram_write_bytes"], launch_ms_under_ncu=avg["ms"],
           dmma_pipe_pct=avg["dmma_pipe_pct"], fp64_pipe_pct=avg["fp64_pipe_pct"], issue_active_pct=avg["issue_active_pct"],
           l2_hit_pct=avg["l2_hit_pct"], shared_conflict_share=avg["shared_conflict_share"], per_launch=rows)
print(json.dumps(out, indent=1))
