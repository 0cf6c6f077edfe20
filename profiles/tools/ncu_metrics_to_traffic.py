#!/usr/bin/env python
"""Turns the metric pass of ncu over one bench step into profiles/traffic.json entries (read by bench.py's roofline).

    ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum,\
gpu__time_duration.sum,... --clock-control none -k regex:'k_extend|k_occlude|k_shade|k_connect|k_generate|k_accumulate' \
        --csv --log-file metrics.csv python bench.py --steps 1 --warmup 1 --only-value
    python profiles/tools/ncu_metrics_to_traffic.py metrics.csv atrium 49766400 width=1920 height=1080 frames_in_flight=24 bounce_limit=8

The capture of `bench.py --steps 1 --warmup 1 --only-value` holds three identical steps (warm-up, timed, the pass with
stage events): the launches of every kernel are split in three and the last third is taken (parts=N overrides). Sums per
kernel over the step: DRAM bytes, warp instructions, thread instructions, ncu's (serialised, cold-cache) time."""
import collections
import csv
import json
import os
import sys


def load(fn):
    with open(fn) as f:
        lines = [ln for ln in f if ln.startswith('"')]
    launches = collections.OrderedDict()
    for row in csv.DictReader(lines):
        d = launches.setdefault(int(row["ID"]), dict(name=row["Kernel Name"].split("(")[0].replace("void ", "").split("<")[0]))
        d[row["Metric Name"]] = float(row["Metric Value"].replace(",", ""))
    return list(launches.values())


def main():
    fn, workload, units = sys.argv[1], sys.argv[2], int(sys.argv[3])
    cfg = {k: int(v) for k, v in (a.split("=") for a in sys.argv[4:] if not a.startswith("--"))}
    parts = cfg.pop("parts", 3)
    L = load(fn)
    per = collections.OrderedDict()
    for l in L:
        per.setdefault(l["name"], []).append(l)
    step = []
    for name, ls in per.items():                      # last 1/parts of each kernel's launches = one step
        n = len(ls) // parts
        step += ls[len(ls) - n:] if n else []
    agg = collections.OrderedDict()
    for l in step:
        a = agg.setdefault(l["name"], dict(launches=0, dram_bytes=0.0, warp_inst=0.0, thread_inst=0.0, time_ms_ncu=0.0))
        a["launches"] += 1
        a["dram_bytes"] += l["dram__bytes_read.sum"] + l["dram__bytes_write.sum"]
        a["warp_inst"] += l["smsp__inst_executed.sum"]
        a["thread_inst"] += l["smsp__thread_inst_executed.sum"]
        a["time_ms_ncu"] += l["gpu__time_duration.sum"] / 1e6
    for a in agg.values():
        a["threads_per_inst"] = round(a["thread_inst"] / max(a["warp_inst"], 1.0), 2)
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "traffic.json")
    try:
        tj = json.load(open(path))
    except (OSError, ValueError):
        tj = {}
    if "unit" in tj and "kernels" not in tj.get("atrium", {"kernels": 1}):
        tj = {}                                     # round-1 format
    entry = dict(config=cfg, kernels=agg, source=os.path.basename(fn))
    entry["rays" if workload == "soup" else "paths"] = units
    if workload == "soup":
        entry["triangles"] = cfg.pop("triangles", None)
        entry["config"] = cfg
    tj[workload] = entry
    tj["note"] = ("per workload: sums over the launches of ONE bench step from ncu's metric pass of the same command "
                  "(dram__bytes_read + dram__bytes_write, smsp__inst_executed, smsp__thread_inst_executed); `paths` / `rays` = units of that step")
    json.dump(tj, open(path, "w"), indent=1)
    print(json.dumps(entry, indent=1))


if __name__ == "__main__":
    main()
