#!/usr/bin/env python3
"""ncu csv (dram__bytes_read.sum, dram__bytes_write.sum, gpu__time_duration.sum for every launch of ONE training step)
-> profiles/r02_traffic.json: DRAM bytes per launch for each kernel family bench.py reports, plus a per-kernel table.

    ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -s <first launch
        of a step> -c <launches per step> --csv --log-file x.csv python bench.py --steps 1 --warmup 3 ...   (AB_NO_GRAPH=1)
    python profiles/traffic_from_ncu.py x.csv f16
"""
import collections
import csv
import json
import os
import re
import sys

FAMILY = [  # kernel name substring -> bench.py family
    ("conv_tc_kernel", "conv_tc"), ("wgrad_tc_kernel", "wgrad_tc"), ("wgrad_tc_reduce", "wgrad_tc"),
    ("bn_", "ew_f16"), ("maxpool", "ew_f16"), ("resize_bilinear", "ew_f16"), ("upsample_zero", "ew_f16"), ("channel_sum", "ew_f16"), ("planes_to_nhwc8", "ew_f16"),
    ("pack_", "pack"), ("loss_", "loss"), ("image_to_nhwc", "input"), ("sgd", "optimizer"), ("adam", "optimizer"),
]
UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1.0, "ms": 1e3}


def main():
    src, prec = sys.argv[1], sys.argv[2]
    per = collections.OrderedDict()  # launch id -> {name, read, write, us}
    for row in csv.reader(open(src, errors="ignore")):
        if len(row) < 10 or not row[0].isdigit():
            continue
        d = per.setdefault(row[0], {"name": re.sub(r"\(.*", "", row[4]).split("::")[-1]})
        val = float(row[-1].replace(",", "")) * UNIT.get(row[-2], 1.0)
        if row[-3].startswith("dram__bytes_read"):
            d["read"] = val
        elif row[-3].startswith("dram__bytes_write"):
            d["write"] = val
        elif row[-3].startswith("gpu__time_duration"):
            d["us"] = val
    steps = max(1, sum(1 for d in per.values() if d["name"].startswith("loss_kernel")))   # one loss kernel per training step in the capture
    fam = collections.defaultdict(lambda: {"launches": 0, "dram_bytes": 0.0, "us": 0.0})
    kern = collections.defaultdict(lambda: {"launches": 0, "dram_bytes": 0.0, "us": 0.0})
    for d in per.values():
        f = next((fm for sub, fm in FAMILY if sub in d["name"]), "other")
        for tbl, key in ((fam, f), (kern, d["name"])):
            tbl[key]["launches"] += 1
            tbl[key]["dram_bytes"] += d.get("read", 0.0) + d.get("write", 0.0)
            tbl[key]["us"] += d.get("us", 0.0)
    out_path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "r02_traffic.json")
    try:
        out = json.load(open(out_path))
    except (OSError, ValueError):
        out = {}
    out["source"] = "ncu dram__bytes_read.sum + dram__bytes_write.sum, every launch of one training step (" + os.path.basename(src) + ")"
    out["steps_in_capture"] = steps
    out[prec] = {k: {"launches_per_step": v["launches"] / steps, "dram_bytes_per_launch": v["dram_bytes"] / v["launches"],
                     "dram_bytes_per_step": v["dram_bytes"] / steps, "us_per_step_under_ncu": v["us"] / steps,
                     "gbs_under_ncu": v["dram_bytes"] / v["us"] / 1e3 if v["us"] else None} for k, v in fam.items()}
    out[prec + "_kernels"] = {k: {"launches_per_step": v["launches"] / steps, "dram_mb_per_step": v["dram_bytes"] / 1e6 / steps, "us_per_step": v["us"] / steps,
                                  "gbs": v["dram_bytes"] / v["us"] / 1e3 if v["us"] else None}
                              for k, v in sorted(kern.items(), key=lambda kv: -kv[1]["us"])}
    json.dump(out, open(out_path, "w"), indent=1)
    for k, v in out[prec + "_kernels"].items():
        print(f"{k[:44]:44s} {v['launches_per_step']:6.1f} {v['dram_mb_per_step']:10.1f} MB {v['us_per_step']:9.1f} us {v['gbs'] or 0:8.1f} GB/s")


if __name__ == "__main__":
    main()
