"""Summaries of ncu output for profiles/ (run in the build container; ncu reads reports without a GPU).
  python scripts/ncu_summary.py launches RAW.csv OUT.csv "header text"     # --metrics gpu__time_duration.sum launch list
  python scripts/ncu_summary.py full REPORT.ncu-rep OUT.csv "header text"  # --set full capture -> selected metrics per launch
"""
import csv, io, subprocess, sys, collections, re

KEEP = ["launch__grid_size", "launch__block_size", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__m_xbar2l1tex_read_bytes.sum",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_shared_mem",
        "launch__waves_per_multiprocessor", "smsp__inst_executed.sum", "sm__cycles_active.avg", "sm__cycles_elapsed.max",
        "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active"]


def launches(raw, out, header):
    rows = [l for l in open(raw) if not l.startswith("==")]
    r = list(csv.reader(io.StringIO("".join(rows))))
    h = r[0]
    ki, vi, ui = h.index("Kernel Name"), h.index("Metric Value"), h.index("Metric Unit")
    agg = collections.OrderedDict()
    for row in r[1:]:
        if len(row) <= vi:
            continue
        name = re.sub(r"\(.*", "", row[ki]).strip()
        name = re.sub(r"<.*", "", name)
        v = float(row[vi].replace(",", ""))
        us = v / 1e3 if row[ui] in ("ns", "nsecond") else (v if row[ui] in ("us", "usecond") else v * 1e3)
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1
        a[1] += us
    tot = sum(a[1] for a in agg.values())
    with open(out, "w") as f:
        f.write(f"# {header}\n# ncu --metrics gpu__time_duration.sum --clock-control none ; per-launch times are cold-cache and "
                f"serialised: compare SHARES\n# launches {sum(a[0] for a in agg.values())}, total {tot/1e3:.2f} ms\n")
        f.write("kernel,launches,total_us,share,avg_us\n")
        for k, (n, us) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write(f"{k},{n},{us:.1f},{us/tot:.4f},{us/n:.2f}\n")


def full(rep, out, header):
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    r = list(csv.reader(io.StringIO(txt)))
    h, units, data = r[0], r[1], r[2:]
    ki = h.index("Kernel Name")
    with open(out, "w") as f:
        f.write(f"# {header}\n")
        f.write("metric,unit," + ",".join(f"launch{i+1}" for i in range(len(data))) + "\n")
        f.write("kernel,," + ",".join('"' + d[ki].replace('pdac::', '') + '"' for d in data) + "\n")
        for m in KEEP:
            if m in h:
                i = h.index(m)
                f.write(f"{m},{units[i]}," + ",".join(d[i].replace(",", "") for d in data) + "\n")


if __name__ == "__main__":
    {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2], sys.argv[3], sys.argv[4])
