#!/usr/bin/env python
"""Turns the ncu outputs that come back in gpurun_out/ into the small, tracked summaries under profiles/.

  python tools/summarize_profiles.py gpurun_out/launches_X.csv gpurun_out/prof_X.ncu-rep r01

Writes profiles/<tag>_launches.csv (kernel, launches, total us, share of one boosting round),
profiles/<tag>_ncu_full.csv (selected raw metrics of every captured launch) and
profiles/<tag>_roofline.json (per-launch DRAM traffic of the histogram kernel, read by bench.py)."""
import collections
import csv
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KEEP = ("gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared_op_atom.sum",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_atom.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "smsp__inst_executed.sum", "smsp__inst_executed_op_shared_atom.sum", "launch__registers_per_thread",
        "launch__grid_size", "launch__block_size", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active")


def launches(path, tag):
    lines = [l for l in open(path) if not l.startswith("==")]
    seq = []
    for r in csv.DictReader(lines):
        name = re.sub(r"\(.*", "", r["Kernel Name"]).replace("void ", "").replace("cb::", "")
        v = float(r["Metric Value"].replace(",", ""))
        v = v / 1e3 if r["Metric Unit"] == "ns" else (v * 1e3 if r["Metric Unit"] == "ms" else v)
        seq.append((name, v))
    starts = [i for i, (n, _) in enumerate(seq) if n.startswith("grad_kernel")]
    if len(starts) < 2:
        raise SystemExit(f"{path}: need at least two boosting rounds in the launch list, found {len(starts)}")
    a, b = (starts[3], starts[4]) if len(starts) > 4 else (starts[-2], starts[-1])
    agg = collections.OrderedDict()
    for n, v in seq[a:b]:
        c = agg.setdefault(n, [0, 0.0])
        c[0] += 1
        c[1] += v
    tot = sum(v for _, v in agg.values())
    out = os.path.join(ROOT, "profiles", f"{tag}_launches.csv")
    with open(out, "w", newline="") as fh:
        w = csv.writer(fh)
        w.writerow(["# one boosting round (the 4th of the run), ncu gpu__time_duration.sum, cold cache, serialised"])
        w.writerow(["kernel", "launches", "total_us", "share"])
        for n, (c, v) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            w.writerow([n, c, round(v, 1), round(v / tot, 4)])
        w.writerow(["ROUND TOTAL", sum(c for c, _ in agg.values()), round(tot, 1), 1.0])
        w.writerow([])
        w.writerow(["# launch sequence of that round (kernel, us)"])
        for n, v in seq[a:b]:
            w.writerow([n, round(v, 1)])
        w.writerow([])
        w.writerow(["# one-off preparation (sketch, binning) and scoring launches"])
        pre = collections.OrderedDict()
        for n, v in seq[:starts[0]] + [x for x in seq[starts[-1]:] if "predict" in x[0] or "bin_rows" in x[0]]:
            c = pre.setdefault(n, [0, 0.0])
            c[0] += 1
            c[1] += v
        for n, (c, v) in pre.items():
            w.writerow([n, c, round(v, 1)])
    print("wrote", out)


def full(rep, tag):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rd = list(csv.reader(raw.splitlines()))
    hdr, units, rows = rd[0], rd[1], rd[2:]
    idx = {h: i for i, h in enumerate(hdr)}
    cols = ["ID", "Kernel Name"] + [k for k in KEEP if k in idx]
    out = os.path.join(ROOT, "profiles", f"{tag}_ncu_full.csv")
    with open(out, "w", newline="") as fh:
        w = csv.writer(fh)
        w.writerow(cols)
        w.writerow(["", ""] + [units[idx[k]] for k in cols[2:]])
        for r in rows:
            w.writerow([r[idx["ID"]], re.sub(r"\(.*", "", r[idx["Kernel Name"]])[:48]] + [r[idx[k]] for k in cols[2:]])
    print("wrote", out)

    def num(r, k):
        v = float(r[idx[k]].replace(",", ""))
        u = units[idx[k]].lower()
        return v * {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}.get(u, 1)
    hist = [r for r in rows if "hist_accum" in r[idx["Kernel Name"]]]
    if hist:
        tr = [num(r, "dram__bytes_read.sum") + num(r, "dram__bytes_write.sum") for r in hist]
        # rows a launch scanned = shared-memory atomic warp instructions / 8 (4 code bytes x (g, h) per row word);
        # algorithmic bytes of the SAME launches by SURVEY 8d's formula R * (F + 8 + 4 [below the root]) with F = 150
        rows_ = [num(r, "smsp__inst_executed_op_shared_atom.sum") / 8 for r in hist]
        alg = [x * (150 + 8 + (0 if "1>" in r[idx["Kernel Name"]].split("(")[0] else 4)) for x, r in zip(rows_, hist)]
        js = {"kernel": "hist_accum_cf_kernel", "launches_captured": len(hist),
              "dram_bytes_per_launch": [int(x) for x in tr], "dram_bytes_per_launch_mean": int(sum(tr) / len(tr)),
              "rows_per_launch": [int(x) for x in rows_], "algorithmic_bytes_per_launch": [int(x) for x in alg],
              "traffic_over_algorithmic": round(sum(tr) / sum(alg), 3),
              "note": "one ncu --set full capture of consecutive launches (levels 0.. of one round); dram__bytes_read.sum + "
                      "dram__bytes_write.sum; rows from the launch's shared-atomic instruction count (layout: 128-byte rows, "
                      "one-hot groups packed), histogram write-out not counted"}
        with open(os.path.join(ROOT, "profiles", f"{tag}_roofline.json"), "w") as fh:
            json.dump(js, fh, indent=1)
        print("hist traffic per launch (MB):", [round(x / 1e6, 1) for x in tr])


if __name__ == "__main__":
    launches(sys.argv[1], sys.argv[3])
    if os.path.exists(sys.argv[2]):
        full(sys.argv[2], sys.argv[3])
