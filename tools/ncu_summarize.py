#!/usr/bin/env python
"""Summarise an ncu CSV launch list (gpu__time_duration.sum [+ dram__bytes_*]) per kernel.

    python tools/ncu_summarize.py launches.csv [--traffic-json out.json --steps-in-capture N]

Writes a markdown table to stdout.  With --traffic-json also writes the DRAM bytes of the
tcgen05 convolution launches per step (what bench.py reports as roofline.traffic)."""
import argparse
import collections
import csv
import json
import re


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("csv")
    ap.add_argument("--traffic-json")
    ap.add_argument("--steps-in-capture", type=float, default=None)
    ap.add_argument("--note", default="")
    a = ap.parse_args()
    rows = []
    with open(a.csv) as f:
        lines = [l for l in f if l.startswith('"')]
    rd = csv.reader(lines)
    hdr = next(rd)
    ix = {h: i for i, h in enumerate(hdr)}
    per = collections.defaultdict(lambda: collections.defaultdict(float))  # (id) -> metric -> value
    name = {}
    for r in rd:
        if len(r) < len(hdr):
            continue
        kid = int(r[ix["ID"]])
        name[kid] = r[ix["Kernel Name"]]
        v = float(r[ix["Metric Value"]].replace(",", ""))
        unit = r[ix["Metric Unit"]]
        m = r[ix["Metric Name"]]
        if m == "gpu__time_duration.sum":
            v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(unit, 1e-3)
        elif m.startswith("dram__bytes"):
            v *= {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1.0)
        per[kid][m] = v
    agg = collections.defaultdict(lambda: [0, 0.0, 0.0])
    for kid, m in per.items():
        short = re.sub(r"\(.*", "", name[kid]).replace("void ", "").replace("ai4e::", "")
        short = re.sub(r"\(int\)", "", short)
        g = agg[short]
        g[0] += 1
        g[1] += m.get("gpu__time_duration.sum", 0.0)
        g[2] += m.get("dram__bytes_read.sum", 0.0) + m.get("dram__bytes_write.sum", 0.0)
    tot = sum(g[1] for g in agg.values())
    print("| kernel | launches | total us | share | DRAM MB |")
    print("|---|---|---|---|---|")
    for k, g in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"| `{k}` | {g[0]} | {g[1]:.1f} | {g[1] / tot:.3f} | {g[2] / 1e6:.1f} |")
    print(f"\ntotal {tot:.1f} us over {sum(g[0] for g in agg.values())} launches")
    if a.traffic_json:
        steps = a.steps_in_capture or 1.0
        conv = [g for k, g in agg.items() if "conv_igemm" in k or "bottleneck_tail" in k or "stem_pool" in k]
        byts = sum(g[2] for g in conv) / steps
        json.dump({"kernel": "conv_igemm_kernel + bottleneck_tail_kernel + stem_pool_kernel", "dram_bytes_per_step": byts, "launches_per_step": sum(g[0] for g in conv) / steps,
                   "us_per_step_under_ncu": sum(g[1] for g in conv) / steps, "steps_in_capture": steps,
                   "note": a.note}, open(a.traffic_json, "w"), indent=1)


if __name__ == "__main__":
    main()
