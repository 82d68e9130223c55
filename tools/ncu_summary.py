"""Turn an ncu launch list (CSV of gpu__time_duration.sum) and one `--set full` report into the markdown summary kept
under profiles/.   python tools/ncu_summary.py <launches.csv> <report.ncu-rep> > profiles/<name>.md"""
import csv, collections, subprocess, sys

def launch_table(path):
    rows = [r for r in csv.reader(open(path)) if len(r) > 5]
    hdr = next(r for r in rows if "Kernel Name" in r)
    ik, iv = hdr.index("Kernel Name"), hdr.index("Metric Value")
    agg = collections.defaultdict(list)
    for r in rows:
        if r is hdr or len(r) <= iv: continue
        try: agg[r[ik]].append(float(r[iv].replace(",", "")))
        except ValueError: pass
    # ncu reports ns (or us): normalise through the unit column when present
    iu = hdr.index("Metric Unit") if "Metric Unit" in hdr else None
    unit = next((r[iu] for r in rows if r is not hdr and iu is not None and len(r) > iu), "ns")
    scale = {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(unit, 1e-3)
    tot = sum(sum(v) for v in agg.values())
    out = ["| kernel | launches | avg us | share |", "|---|---|---|---|"]
    for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
        out.append("| `%s` | %d | %.1f | %.3f |" % (k[:70], len(v), sum(v) / len(v) * scale, sum(v) / tot))
    return "\n".join(out), sum(len(v) for v in agg.values())

WANT = ["dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "gpu__time_duration.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "launch__registers_per_thread",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active"]

def raw_table(rep):
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(txt.splitlines()))
    hdr, units, vals = rows[0], rows[1], rows[2]
    out = ["| metric | unit | value |", "|---|---|---|"]
    for i, h in enumerate(hdr):
        if h in WANT or (h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio")):
            try:
                if h.startswith("smsp__average") and float(vals[i].replace(",", "")) < 0.15: continue
            except ValueError: pass
            out.append("| %s | %s | %s |" % (h, units[i], vals[i]))
    return "\n".join(out)

if __name__ == "__main__":
    t, n = launch_table(sys.argv[1])
    print("## Launch list (%d launches, cold-cache serialised times: compare shares)\n\n%s\n" % (n, t))
    print("## K2 raw metrics (per launch)\n\n%s" % raw_table(sys.argv[2]))
