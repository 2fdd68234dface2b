"""Round-2 ncu post-processing (build container; no GPU needed):
  python scripts/ncu_r02.py gpurun_out/r2c_launches_raw.csv gpurun_out/r2c_full_raw.csv profiles/ r02
-> profiles/r02_launches_summary.csv   every kernel of ONE smp-Unet train step at batch 64 (third step of one_step.py):
                                       launches, device time, share, DRAM bytes read / written
   profiles/r02_full_summary.csv       the --set full capture: one row per launch of conv_halo / wgrad_halo / bn_bwd_apply
   profiles/r02_traffic.json           mean DRAM bytes per launch of those kernels (what bench.py reports as `traffic`)"""
import collections, csv, io, json, re, sys

raw, full, outdir, tag = sys.argv[1:5]


def short(name):
    name = name.replace("pdac::", "")
    name = re.sub(r"\(.*", "", name).strip()
    return re.sub(r"^void ", "", name)


# ---- launch list: rows are (ID, kernel, metric, value); keep the launches of the last step
rows = [l for l in open(raw) if not l.startswith("==")]
r = list(csv.reader(io.StringIO("".join(rows))))
h = r[0]
ii, ki, mi, vi = h.index("ID"), h.index("Kernel Name"), h.index("Metric Name"), h.index("Metric Value")
launch = collections.OrderedDict()
for row in r[1:]:
    if len(row) <= vi:
        continue
    d = launch.setdefault(int(row[ii]), {"k": short(row[ki])})
    d[row[mi]] = float(row[vi].replace(",", ""))
ids = sorted(launch)
# the three steps launch the same kernels except for first-step one-time work: find the step length from the conv count
conv_ids = [i for i in ids if launch[i]["k"].startswith("conv_halo_kernel")]
per_step = len(conv_ids) // 3
first_of_last = conv_ids[2 * per_step]
# the step starts a few launches before its first conv (input prep, weight packing of the previous step belongs to it)
prep = [i for i in ids if i < first_of_last and launch[i]["k"].startswith("stem_prep")]
start = prep[-1] if prep else first_of_last
step = [launch[i] for i in ids if i >= start]
agg = collections.OrderedDict()
for d in step:
    name = re.sub(r"<.*", "", d["k"])
    a = agg.setdefault(name, [0, 0.0, 0.0, 0.0])
    a[0] += 1
    a[1] += d.get("gpu__time_duration.sum", 0.0) / 1e3
    a[2] += d.get("dram__bytes_read.sum", 0.0)
    a[3] += d.get("dram__bytes_write.sum", 0.0)
tot = sum(a[1] for a in agg.values())
with open(f"{outdir}/{tag}_launches_summary.csv", "w") as f:
    f.write(f"# {tag}: ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none ; "
            f"python scripts/one_step.py 64 3 (smp Unet-ResNet18 train step, 512x512, batch 64, eager, deterministic reductions)\n"
            f"# kernels of the THIRD step ({sum(a[0] for a in agg.values())} launches, {tot / 1e3:.2f} ms serialised, cold-cache per launch: compare SHARES)\n")
    f.write("kernel,launches,total_us,share,avg_us,dram_read_MB,dram_write_MB\n")
    for k, (n, us, rd, wr) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        f.write(f"{k},{n},{us:.1f},{us / tot:.4f},{us / n:.2f},{rd / 1e6:.1f},{wr / 1e6:.1f}\n")

# ---- full capture
fr = list(csv.reader(open(full)))
fh, fu, fd = fr[0], fr[1], fr[2:]
KEEP = ["Kernel Name", "launch__grid_size", "launch__block_size", "gpu__time_duration.sum", "dram__bytes_read.sum",
        "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "launch__waves_per_multiprocessor"]
cols = [fh.index(k) for k in KEEP if k in fh]


def num(d, key):
    i = fh.index(key)
    v = float(d[i].replace(",", ""))
    u = fu[i]
    scale = {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1.0, "us": 1e-6, "ns": 1e-9, "ms": 1e-3, "usecond": 1e-6,
             "nsecond": 1e-9, "msecond": 1e-3, "second": 1.0}.get(u, 1.0)
    return v * scale


with open(f"{outdir}/{tag}_full_summary.csv", "w") as f:
    f.write(f"# {tag}: ncu --set full --clock-control none -k regex:conv_halo_kernel|wgrad_halo_kernel|bn_bwd_apply_kernel -s 214 -c 107 ; "
            "python scripts/one_step.py 64 3 : every launch of the three kernels in the third smp-Unet train step at batch 64 "
            "(the bench configuration), in launch order\n")
    f.write(",".join(fh[c] for c in cols) + "\n")
    f.write(",".join(fu[c] for c in cols) + "\n")
    for d in fd:
        f.write(",".join('"' + short(d[c]) + '"' if fh[c] == "Kernel Name" else d[c].replace(",", "") for c in cols) + "\n")

traffic = {}
for d in fd:
    name = re.sub(r"<.*", "", short(d[fh.index("Kernel Name")]))
    t = traffic.setdefault(name, {"launches": 0, "dram_bytes": 0.0, "seconds": 0.0, "tensor_pct": [], "warps_pct": []})
    t["launches"] += 1
    t["dram_bytes"] += num(d, "dram__bytes_read.sum") + num(d, "dram__bytes_write.sum")
    t["seconds"] += num(d, "gpu__time_duration.sum")
    t["tensor_pct"].append(float(d[fh.index("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active")]))
    t["warps_pct"].append(float(d[fh.index("sm__warps_active.avg.pct_of_peak_sustained_active")]))
out = {}
for k, t in traffic.items():
    out[k] = {"dram_bytes_per_launch": t["dram_bytes"] / t["launches"], "launches": t["launches"],
              "dram_gbs_under_ncu": t["dram_bytes"] / t["seconds"] / 1e9,
              "tensor_pipe_active_pct_mean": sum(t["tensor_pct"]) / len(t["tensor_pct"]),
              "tensor_pipe_active_pct_max": max(t["tensor_pct"]),
              "warps_active_pct_mean": sum(t["warps_pct"]) / len(t["warps_pct"]),
              "note": f"mean over ALL {t['launches']} launches of this kernel in one smp-Unet train step at batch 64 "
                      f"(profiles/{tag}_full_summary.csv), the same set bench.py's `achieved` averages over"}
json.dump(out, open(f"{outdir}/{tag}_traffic.json", "w"), indent=1)
print(json.dumps(out, indent=1))
