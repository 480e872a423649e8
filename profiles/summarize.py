#!/usr/bin/env python
"""Turn ncu output brought back in gpurun_out/ into the tracked summaries under profiles/.

  python profiles/summarize.py launches <launches.csv> <out.md>       per-kernel launch list (gpu__time_duration.sum)
  python profiles/summarize.py raw <raw.csv> <out.md> [key=regex ...]   the same from `ncu -i rep --page raw --csv` exported on the GPU box
  python profiles/summarize.py stalls <source.csv> <out.md>             warp-state samples by reason / by opcode + the hottest non-FFMA2
                                                                        instructions, from `ncu -i rep --page source --csv`
  python profiles/summarize.py full <report.ncu-rep> <out.md> [key=regex ...]
        one block of headline metrics per profiled launch; `key=regex` pairs also record
        dram read+write bytes per launch of the first kernel matching regex into profiles/traffic.json
        (bench.py reports it as roofline.traffic)
Runs here (no GPU): it only reads the report with `ncu -i`.
"""
import collections
import csv
import io
import json
import os
import re
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))

METRICS = [
    ("gpu__time_duration.sum", "duration"),
    ("launch__grid_size", "grid"),
    ("launch__block_size", "block"),
    ("launch__registers_per_thread", "regs/thread"),
    ("launch__occupancy_limit_registers", "occupancy limit (regs), CTAs/SM"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "achieved occupancy %"),
    ("dram__bytes_read.sum", "DRAM read"),
    ("dram__bytes_write.sum", "DRAM write"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM throughput % of peak"),
    ("lts__t_sector_hit_rate.pct", "L2 hit rate %"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM throughput % of peak"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue slots busy %"),
    ("sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "FMA pipe active %"),
    ("sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "FP64 pipe active %"),
    ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor pipe active %"),
    ("sm__inst_executed_pipe_tensor_op_dmma.avg.pct_of_peak_sustained_active", "DMMA pipe %"),
    ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "shared-memory bank conflicts"),
]

BYTES = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}


def launches(csv_path, out_path):
    rows = [r for r in csv.reader(open(csv_path)) if len(r) > 10]
    hdr = rows[0]
    ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    agg = collections.OrderedDict()
    for r in rows[1:]:
        v = float(r[vi].replace(",", ""))
        if r[ui] == "us":
            v *= 1e3
        elif r[ui] == "ms":
            v *= 1e6
        a = agg.setdefault(r[ki], [0, 0.0])
        a[0] += 1
        a[1] += v
    total = sum(v[1] for v in agg.values())
    with open(out_path, "w") as f:
        f.write("| launches | total ms | share | kernel |\n|---:|---:|---:|---|\n")
        for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write("| %d | %.3f | %.1f%% | `%s` |\n" % (v[0], v[1] / 1e6, 100 * v[1] / total, k[:140]))
        f.write("\n%d launches, %.1f ms of device time (ncu-serialised, cold cache: compare shares, not absolutes)\n" % (
            sum(v[0] for v in agg.values()), total / 1e6))


def stalls(src_csv, out_path):
    rows = list(csv.reader(open(src_csv)))
    kernel = rows[0][1] if len(rows[0]) > 1 else ""
    hdr, data = rows[1], rows[2:]
    ix = {h: i for i, h in enumerate(hdr)}
    cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
    by_reason, by_op, lds = collections.Counter(), collections.Counter(), collections.Counter()
    total, lines = 0, []
    for r in data:
        n = int(r[ix["# Samples"]] or 0)
        total += n
        src = r[ix["Source"]].strip()
        op = (src.split()[1] if src.startswith("@") else src.split()[0]) if src else ""
        by_op[op.split(".")[0]] += n
        st = {c[6:]: int(r[ix[c]] or 0) for c in cols if int(r[ix[c]] or 0)}
        for k, v in st.items():
            by_reason[k] += v
        if op.startswith("LDS"):
            lds[op + " wavefronts"] += int(r[ix["L1 Wavefronts Shared"]] or 0)
            lds[op + " instructions"] += int(r[ix["Instructions Executed"]] or 0)
        lines.append((n, src[:70], st))
    with open(out_path, "w") as f:
        f.write("Source: `%s` (ncu --set full --import-source on, source page)\n\nKernel: `%s`\n\n" % (os.path.basename(src_csv), kernel[:200]))
        f.write("%d warp-state samples.\n\n| stall reason | samples | share |\n|---|---:|---:|\n" % total)
        for k, v in by_reason.most_common(10):
            f.write("| %s | %d | %.1f%% |\n" % (k, v, 100.0 * v / max(total, 1)))
        f.write("\n| opcode | samples | share |\n|---|---:|---:|\n")
        for k, v in by_op.most_common(8):
            f.write("| %s | %d | %.1f%% |\n" % (k, v, 100.0 * v / max(total, 1)))
        for op in sorted({k.rsplit(" ", 1)[0] for k in lds}):
            if lds[op + " instructions"]:
                f.write("\n`%s`: %.2f shared-memory wavefronts per instruction (%d instructions)\n" % (op, lds[op + " wavefronts"] / lds[op + " instructions"], lds[op + " instructions"]))
        f.write("\nHottest instructions that are not FFMA2:\n\n| samples | instruction | stall reasons |\n|---:|---|---|\n")
        for n, src, st in sorted([l for l in lines if not l[1].startswith("FFMA2")], key=lambda l: -l[0])[:12]:
            f.write("| %d | `%s` | %s |\n" % (n, src, ", ".join("%s %d" % kv for kv in sorted(st.items(), key=lambda kv: -kv[1])[:3])))


def full(rep_path, out_path, pairs, raw_csv=None):
    raw = open(raw_csv).read() if raw_csv else subprocess.run(["ncu", "-i", rep_path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    col = {h: i for i, h in enumerate(hdr)}
    traffic_path = os.path.join(HERE, "traffic.json")
    try:
        traffic = json.load(open(traffic_path))
    except Exception:
        traffic = {}
    with open(out_path, "w") as f:
        f.write("Source: `%s` (ncu --set full --clock-control none --import-source on)\n\n" % os.path.basename(rep_path))
        for r in rows[2:]:
            name = r[col["Kernel Name"]]
            f.write("### `%s`\n\n| metric | value |\n|---|---|\n" % name[:160])
            for m, label in METRICS:
                if m in col:
                    f.write("| %s (`%s`) | %s %s |\n" % (label, m, r[col[m]], units[col[m]]))
            rd = float(r[col["dram__bytes_read.sum"]].replace(",", "")) * BYTES.get(units[col["dram__bytes_read.sum"]], 1.0)
            wr = float(r[col["dram__bytes_write.sum"]].replace(",", "")) * BYTES.get(units[col["dram__bytes_write.sum"]], 1.0)
            f.write("| DRAM traffic per launch | %.4g bytes |\n\n" % (rd + wr))
            for pair in pairs:
                key, rx = pair.split("=", 1)
                if re.search(rx, name) and key not in traffic.get("_done_" + os.path.basename(rep_path), []):
                    traffic[key] = rd + wr
                    traffic.setdefault("_done_" + os.path.basename(rep_path), []).append(key)
    for k in [k for k in traffic if k.startswith("_done_")]:
        del traffic[k]
    json.dump(traffic, open(traffic_path, "w"), indent=1, sort_keys=True)


def traffic(csv_path, key, steps):
    """Sum dram read+write bytes over every launch in an `ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum --csv`
    log that covers `steps` steps of the hot path; stores bytes per step under `key` in profiles/traffic.json."""
    rows = [r for r in csv.reader(open(csv_path)) if len(r) > 10 and r[0].isdigit()]
    total = sum(float(r[-1].replace(",", "")) for r in rows if r[-3].startswith("dram__bytes_"))
    n_launch = len({r[0] for r in rows})
    path = os.path.join(HERE, "traffic.json")
    try:
        t = json.load(open(path))
    except Exception:
        t = {}
    t[key] = total / steps
    t[key + "_launches_per_step"] = n_launch / steps
    json.dump(t, open(path, "w"), indent=1, sort_keys=True)
    print(key, "%.4g bytes per step over %d launches" % (total / steps, n_launch / steps))


if __name__ == "__main__":
    if sys.argv[1] == "traffic":
        traffic(sys.argv[2], sys.argv[3], int(sys.argv[4]) if len(sys.argv) > 4 else 1)
    elif sys.argv[1] == "launches":
        launches(sys.argv[2], sys.argv[3])
    elif sys.argv[1] == "raw":
        full(sys.argv[2], sys.argv[3], sys.argv[4:], raw_csv=sys.argv[2])
    elif sys.argv[1] == "stalls":
        stalls(sys.argv[2], sys.argv[3])
    else:
        full(sys.argv[2], sys.argv[3], sys.argv[4:])
