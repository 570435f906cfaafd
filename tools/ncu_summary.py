"""Per-kernel summary of an `ncu --csv --metrics ...` log (long format: one row per launch x metric).

  python tools/ncu_summary.py gpurun_out/step_metrics.csv [--peak-gbs 6550.7] > profiles/rN_step_kernels.txt
Columns: launches, total ms (cold, serialised: compare shares), share, mean us, DRAM GB/s = (read+write)/time and its
fraction of the measured copy peak, tensor-pipe active %, ncu's own dram/sm throughput % of peak."""
import argparse
import collections
import csv
import io
import json
import os
import re


def short(name):
    name = re.sub(r"\(.*$", "", name)
    name = re.sub(r"^void ", "", name).replace("mft::", "").replace("(anonymous namespace)::", "")
    return name[:58]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("csv")
    ap.add_argument("--peak-gbs", type=float, default=None)
    a = ap.parse_args()
    peak = a.peak_gbs
    if peak is None:
        p = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
        peak = json.load(open(p))["hbm_gbs"] if os.path.exists(p) else 6550.7
    text = open(a.csv).read()
    text = text[text.index('"ID","Process ID"'):]
    per = collections.defaultdict(dict)
    names = {}
    for row in csv.DictReader(io.StringIO(text)):
        i = int(row["ID"])
        names[i] = row["Kernel Name"]
        v = row["Metric Value"].replace(",", "")
        try:
            v = float(v)
        except ValueError:
            continue
        unit = row["Metric Unit"]
        m = row["Metric Name"]
        if m == "gpu__time_duration.sum":
            v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(unit, 1.0)          # -> us
        if m.endswith("bytes_read.sum") or m.endswith("bytes_write.sum") or m == "lts__t_bytes.sum":
            v *= {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1.0)
        per[i][m] = v
    agg = collections.defaultdict(lambda: collections.defaultdict(float))
    for i, m in per.items():
        k = short(names[i])
        g = agg[k]
        t = m.get("gpu__time_duration.sum", 0.0)
        g["n"] += 1
        g["us"] += t
        g["bytes"] += m.get("dram__bytes_read.sum", 0.0) + m.get("dram__bytes_write.sum", 0.0)
        g["l2bytes"] += m.get("lts__t_bytes.sum", 0.0)
        for key, src in (("tensor", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"),
                         ("dram_pct", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
                         ("sm_pct", "sm__throughput.avg.pct_of_peak_sustained_elapsed")):
            g[key] += m.get(src, 0.0) * t                                              # time-weighted
    total = sum(g["us"] for g in agg.values())
    print(f"# {a.csv}: {len(per)} launches, {total/1e3:.3f} ms summed kernel time (cold-cache, serialised); HBM peak {peak:.1f} GB/s (measured copy)")
    print(f"{'kernel':58s} {'n':>4s} {'ms':>8s} {'share':>6s} {'us/launch':>9s} {'DRAM GB/s':>9s} {'of peak':>7s} {'L2 GB/s':>8s} {'tensor%':>7s} {'dram%':>6s} {'sm%':>5s}")
    for k, g in sorted(agg.items(), key=lambda kv: -kv[1]["us"]):
        us = g["us"]
        gbs = g["bytes"] / us / 1e3 if us else 0.0
        l2 = g["l2bytes"] / us / 1e3 if us else 0.0
        print(f"{k:58s} {int(g['n']):4d} {us/1e3:8.3f} {us/total:6.1%} {us/g['n']:9.1f} {gbs:9.0f} {gbs/peak:7.1%} {l2:8.0f} {g['tensor']/us:7.1f} {g['dram_pct']/us:6.1f} {g['sm_pct']/us:5.1f}")


if __name__ == "__main__":
    main()
