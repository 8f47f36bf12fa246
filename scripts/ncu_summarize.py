"""Turn the ncu outputs brought back in gpurun_out/ into the tracked summaries under profiles/.

  python scripts/ncu_summarize.py <launches.csv> <full.ncu-rep | its --page raw --csv export> <tag>
writes profiles/<tag>_ncu_launches.csv (per-launch gpu__time_duration of one bench.py run), profiles/<tag>_ncu_step_share.txt
(per-kernel totals and shares), profiles/<tag>_ncu_full_top_kernels.csv (selected --set full metrics per captured launch) and
profiles/<tag>_ncu_traffic.json (per kernel name as bench.py reports it: launches captured in ONE training step, mean DRAM
read+write bytes per launch, mean duration) -- bench.py sets `roofline.traffic` from it."""
import csv, collections, json, os, re, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
launches, rep, tag = sys.argv[1], sys.argv[2], sys.argv[3]
prof = os.path.join(ROOT, "profiles")

def short(name):
    n = name.split("(")[0].replace("void ", "").replace("<unnamed>::", "")
    return n

rows = []
with open(launches) as f:
    lines = [l for l in f if l.startswith('"')]
for r in csv.DictReader(lines):
    rows.append((short(r["Kernel Name"]), r["Grid Size"], r["Block Size"], float(r["Metric Value"].replace(",", ""))))
with open(os.path.join(prof, tag + "_ncu_launches.csv"), "w") as f:
    f.write("kernel,grid,block,gpu_time_ns\n")
    for k, g, b, v in rows:
        f.write('"%s","%s","%s",%.0f\n' % (k, g, b, v))
agg = collections.defaultdict(lambda: [0, 0.0])
for k, g, b, v in rows:
    agg[k][0] += 1; agg[k][1] += v
tot = sum(v for _, v in agg.values())
with open(os.path.join(prof, tag + "_ncu_step_share.txt"), "w") as f:
    f.write("ncu --metrics gpu__time_duration.sum --clock-control none: %d launches of one bench.py run (set-up + warm-up + timed steps)\n" % len(rows))
    f.write("%-44s %6s %12s %10s %7s\n" % ("kernel", "count", "total_us", "avg_us", "share"))
    for k, (c, v) in sorted(agg.items(), key=lambda x: -x[1][1]):
        f.write("%-44s %6d %12.1f %10.2f %6.1f%%\n" % (k[:44], c, v / 1e3, v / 1e3 / c, 100 * v / tot))

raw = open(rep).read() if rep.endswith(".csv") else subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rr = list(csv.reader(raw.splitlines()))
hdr, data = rr[0], rr[2:]
want = ["Kernel Name", "Grid Size", "Block Size", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__compute_memory_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sector_hit_rate.pct", "l1tex__throughput.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic", "sm__inst_executed_pipe_tensor_op_gmma.sum",
        "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active", "lts__t_bytes.sum", "l1tex__m_xbar2l1tex_read_bytes.sum"]
idx = [(w, hdr.index(w)) for w in want if w in hdr]
units = rr[1]
traffic = collections.defaultdict(list)
with open(os.path.join(prof, tag + "_ncu_full_top_kernels.csv"), "w") as f:
    f.write(",".join('"%s [%s]"' % (w, units[i]) for w, i in idx) + "\n")
    for r in data:
        vals = []
        for w, i in idx:
            vals.append(short(r[i]) if w == "Kernel Name" else r[i])
        f.write(",".join('"%s"' % v for v in vals) + "\n")
        name = short(r[hdr.index("Kernel Name")])
        name = re.sub(r"\(int\)|\(bool\)", "", name)
        m = re.match(r"(conv3x3_\w+)<(\d+)", name)
        m9 = re.match(r"wgrad9_kernel<(\d+)", name)
        key = "%s_bn%s" % (m.group(1), m.group(2)) if m else ("wgrad9_n" + m9.group(1) if m9 else name)
        if key.startswith("wgradm_kernel"):
            key = "wgradm"
        key = re.sub(r"<1>$", "<true>", re.sub(r"<0>$", "<false>", key)) if key.startswith(("loss_grid_kernel", "adam_kernel")) else key
        def num(col):
            i = hdr.index(col); v = float(r[i].replace(",", "")); u = units[i]
            return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
        traffic[key].append({"grid": r[hdr.index("Grid Size")], "dram_bytes": num("dram__bytes_read.sum") + num("dram__bytes_write.sum"),
                             "us": float(r[hdr.index("gpu__time_duration.sum")].replace(",", "")) * (1e-3 if units[hdr.index("gpu__time_duration.sum")] in ("ns", "nsecond") else 1)})
summary = {k: {"launches_in_step": len(v), "dram_bytes_per_launch": sum(e["dram_bytes"] for e in v) / len(v),
               "us_per_launch": sum(e["us"] for e in v) / len(v), "launches": v} for k, v in traffic.items()}
json.dump(summary, open(os.path.join(prof, tag + "_ncu_traffic.json"), "w"), indent=1)
print("kernels:", sorted(traffic))
