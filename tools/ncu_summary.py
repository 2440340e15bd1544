"""Turns gpurun_out/launches_<tag>.csv (+ optional prof_<tag>.ncu-rep) into profiles/<tag>_*.{csv,md} summaries."""
import collections
import csv
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1]
cmd = sys.argv[2] if len(sys.argv) > 2 else "tools/profile_step.py"
out_md = os.path.join(ROOT, "profiles", tag + "_summary.md")
lines = ["# ncu summary %s" % tag, ""]

lcsv = os.path.join(ROOT, "gpurun_out", "launches_%s.csv" % tag)
if os.path.exists(lcsv):
    rows = [r for r in csv.reader(open(lcsv)) if len(r) > 5]
    hdr = rows[0]
    ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    agg = collections.OrderedDict()
    for r in rows[1:]:
        name = r[ki].split("(")[0].replace("thcm::<unnamed>::", "")
        v = float(r[vi].replace(",", ""))
        v = v / 1e3 if r[ui] == "ns" else (v * 1e3 if r[ui] == "ms" else v)
        agg.setdefault(name, []).append(v)
    tot = sum(sum(v) for v in agg.values())
    lines += ["## launch list (`ncu --metrics gpu__time_duration.sum --clock-control none -c 400`, %s)" % cmd, "",
              "cold-cache, serialised launches: compare shares, not absolutes", "",
              "| kernel | launches | avg us | share |", "|---|---|---|---|"]
    for k, v in agg.items():
        lines.append("| %s | %d | %.1f | %.1f%% |" % (k[-60:], len(v), sum(v) / len(v), 100 * sum(v) / tot))
    lines.append("")
    os.makedirs(os.path.join(ROOT, "profiles"), exist_ok=True)
    with open(os.path.join(ROOT, "profiles", "launches_%s.csv" % tag), "w") as f:
        f.write(open(lcsv).read())

rep = os.path.join(ROOT, "gpurun_out", "prof_%s.ncu-rep" % tag)
if os.path.exists(rep):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
            "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
            "smsp__issue_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "smsp__inst_executed.sum",
            "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "lts__t_sectors_op_read.sum", "launch__occupancy_limit_registers",
            "launch__occupancy_limit_shared_mem", "launch__grid_size", "launch__block_size"]
    idx = [(w, hdr.index(w)) for w in want if w in hdr]
    ki = hdr.index("Kernel Name")
    lines += ["## full-set capture (`ncu --set full --clock-control none --import-source on`), one row per captured launch", "",
              "| kernel | " + " | ".join("%s [%s]" % (w.split(".")[0].replace("__", ":"), units[i]) for w, i in idx) + " |",
              "|---|" + "---|" * len(idx)]
    for r in rows[2:]:
        lines.append("| %s | " % r[ki].split("(")[0].replace("thcm::<unnamed>::", "")[-40:] + " | ".join(r[i] for _, i in idx) + " |")
    lines.append("")
    # stall breakdown of each distinct kernel
    seen = set()
    for r in rows[2:]:
        kn = r[ki].split("(")[0].split("::")[-1].split("<")[0]
        if kn in seen:
            continue
        seen.add(kn)
        src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + kn], capture_output=True, text=True).stdout
        srows = list(csv.reader(src.splitlines()))
        if len(srows) < 3:
            continue
        sh = srows[1]
        st = [i for i, h in enumerate(sh) if h.startswith("stall_") and "Not Issued" not in h]
        tot = collections.Counter()
        ninst = 0
        for q in srows[2:]:
            if len(q) < len(sh):
                continue
            for i in st:
                try:
                    tot[sh[i]] += int(q[i])
                except ValueError:
                    pass
            ninst += 1
        s = sum(tot.values()) or 1
        lines += ["### %s: warp-stall samples (first captured launch), %d SASS instructions" % (kn, ninst), "",
                  ", ".join("%s %.1f%%" % (k, 100.0 * v / s) for k, v in tot.most_common(8)), ""]
open(out_md, "w").write("\n".join(lines) + "\n")
print("wrote", out_md)
