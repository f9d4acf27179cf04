#!/usr/bin/env python
"""Turn `ncu -i X.ncu-rep --page raw --csv` exports into the committed artefacts of a round:

  python tools/ncu_tables.py traffic gpurun_out/r2_iter_128_raw.csv 128 30 SH1 480 640 > profiles/r2_traffic.json
  python tools/ncu_tables.py table   gpurun_out/r2_iter_512_raw.csv                     (markdown table on stdout)
  python tools/ncu_tables.py compact gpurun_out/r2_iter_128_raw.csv > profiles/r2_iter_128_ncu.csv   (the columns quoted)
  python tools/ncu_tables.py launches gpurun_out/r2_launches_bench_default.csv          (per-kernel totals of a launch list)
"""
import collections
import csv
import json
import re
import sys

MODES = {0: "ENERGY", 1: "INIT_ALBEDO", 2: "LED_INIT", 3: "PROBE", 4: "ALBEDO", 5: "LIGHT", 6: "POSE", 7: "DIST"}
MODELS = {0: "SH1", 1: "SH2", 2: "LED"}
COLS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "l1tex__m_xbar2l1tex_read_bytes.sum",
        "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct", "launch__registers_per_thread", "launch__grid_size",
        "launch__block_size", "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active"]
SCALE = {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1.0, "us": 1.0, "ms": 1e3, "ns": 1e-3, "ms ": 1e3}


def short(name):
    m = re.search(r"k_sweep<(\d+), (\d+)>", name)
    if m:
        return "k_sweep<%s,%s>" % (MODELS[int(m.group(1))], MODES[int(m.group(2))])
    m = re.search(r"(k_\w+)(<[^>]*>)?", name)
    return (m.group(1) + (m.group(2) or "")) if m else name


def bench_key(name):
    m = re.search(r"k_sweep<(\d+), (\d+)>", name)
    if m:
        return "k_sweep<%s>" % MODES[int(m.group(2))]
    if "k_dist_solve" in name or "k_dist_pcg" in name or "k_dist_block" in name:
        return "k_dist_pcg"
    return short(name)


def load_raw(path):
    rows = list(csv.reader(open(path, errors="replace")))
    hdr, units = rows[0], rows[1]
    ci = {h: i for i, h in enumerate(hdr)}
    out = []
    for r in rows[2:]:
        d = {"name": r[ci["Kernel Name"]]}
        for c in COLS:
            if c in ci and r[ci[c]] != "":
                d[c] = float(r[ci[c]].replace(",", "")) * SCALE.get(units[ci[c]], 1.0)
        out.append(d)
    return out


def main():
    what, path = sys.argv[1], sys.argv[2]
    if what == "launches":
        rows = list(csv.reader(open(path, errors="replace")))
        hi = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
        ci = {h: i for i, h in enumerate(rows[hi])}
        agg = collections.OrderedDict()
        for r in rows[hi + 1:]:
            if len(r) < len(rows[hi]):
                continue
            v = float(r[ci["Metric Value"]]) * SCALE.get(r[ci["Metric Unit"]], 1.0)
            a = agg.setdefault(short(r[ci["Kernel Name"]]), [0, 0.0])
            a[0] += 1; a[1] += v
        tot = sum(v[1] for v in agg.values())
        print("| kernel | launches | total µs | share |\n|---|---:|---:|---:|")
        for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            print("| `%s` | %d | %.1f | %.1f %% |" % (k, v[0], v[1], 100 * v[1] / tot))
        print("\n%d launches, %.1f µs of kernels" % (sum(v[0] for v in agg.values()), tot))
        return
    ks = load_raw(path)
    if what == "traffic":
        G, F, model, H, W = int(sys.argv[3]), int(sys.argv[4]), sys.argv[5], int(sys.argv[6]), int(sys.argv[7])
        out = {"workload": {"grid": G, "frames": F, "model": model, "height": H, "width": W},
               "source": "ncu --set full --clock-control none --import-source on, one steady-state LM iteration (tools/ncu_r2.sh; "
                         "text export of the capture: profiles/%s)" % path.split("/")[-1].replace("_raw.csv", "_ncu.csv"),
               "kernels": {}}
        for k in ks:
            out["kernels"][bench_key(k["name"])] = {"kernel": short(k["name"]), "dram_read_bytes": k["dram__bytes_read.sum"],
                                                    "dram_write_bytes": k["dram__bytes_write.sum"],
                                                    "duration_us_under_ncu": k["gpu__time_duration.sum"]}
        json.dump(out, sys.stdout, indent=1)
        print()
    elif what == "compact":
        w = csv.writer(sys.stdout)
        w.writerow(["kernel"] + COLS)
        for k in ks:
            w.writerow([short(k["name"])] + [k.get(c, "") for c in COLS])
    elif what == "table":
        print("| kernel | µs (ncu) | DRAM read MB | DRAM write MB | L2→L1 MB | warp insts M | issue active % | warps active % | SM thr % | regs | grid × block |")
        print("|---|---:|---:|---:|---:|---:|---:|---:|---:|---:|---|")
        for k in ks:
            print("| `%s` | %.1f | %.2f | %.2f | %.1f | %.2f | %.0f | %.0f | %.0f | %d | %d × %d |" % (
                short(k["name"]), k["gpu__time_duration.sum"], k["dram__bytes_read.sum"] / 1e6, k["dram__bytes_write.sum"] / 1e6,
                k.get("l1tex__m_xbar2l1tex_read_bytes.sum", 0) / 1e6, k["smsp__inst_executed.sum"] / 1e6,
                k["smsp__issue_active.avg.pct_of_peak_sustained_active"], k["sm__warps_active.avg.pct_of_peak_sustained_active"],
                k["sm__throughput.avg.pct_of_peak_sustained_elapsed"], k["launch__registers_per_thread"], k["launch__grid_size"],
                k["launch__block_size"]))


if __name__ == "__main__":
    main()
