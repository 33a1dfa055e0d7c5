#!/usr/bin/env python3
"""Summarise result folders written by main.py (this framework's or the reference's).

    python benchmarks/summarize_results.py results/experiment2_8_job [more folders] [--skip 5] [--json out.json]

Same aggregation as BASELINE.md: per experiment prefix, mean over batches >= --skip of every phase
column, then the max over workers; images/sec = workers * data_len / batch_ms.  Sub-folders
(fusion sweeps: one per bucket size in KiB) are reported separately.
"""
from __future__ import annotations

import argparse
import csv
import glob
import json
import os
import re
from collections import defaultdict


def summarise_folder(folder: str, skip: int):
    per = defaultdict(list)
    for path in sorted(glob.glob(os.path.join(folder, "*_times.csv"))):
        m = re.match(r"(.+)_(\d+)_(\d+)_times\.csv$", os.path.basename(path))
        if not m:
            continue
        with open(path) as f:
            rows = list(csv.reader(f, skipinitialspace=True))
        if len(rows) < 2:
            continue
        header, body = rows[0], rows[1:]
        body = body[skip:] if len(body) > skip else body
        means = {}
        for i, h in enumerate(header):
            if h == "experiment_name":
                continue
            vals = [float(r[i]) for r in body if r[i] not in ("", "None")]
            means[h] = sum(vals) / len(vals) if vals else 0.0
        per[m.group(1)].append(means)
    out = {}
    for name, workers in per.items():
        cols = workers[0].keys()
        mx = {c: max(w.get(c, 0.0) for w in workers) for c in cols}
        n = len(workers)
        batch_ms = mx.get("batch", 0.0)
        out[name] = {"workers": n, "batch_ms": batch_ms,
                     "images_per_sec": (n * mx.get("data_len", 0.0) / batch_ms * 1000.0) if batch_ms else 0.0,
                     "phases_ms": {c: mx[c] for c in cols if c not in ("batch", "batch_count", "data_len")}}
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("folders", nargs="+")
    ap.add_argument("--skip", type=int, default=5)
    ap.add_argument("--json", default=None)
    args = ap.parse_args()
    report = {}
    for folder in args.folders:
        subs = [folder] + sorted((d for d in glob.glob(os.path.join(folder, "*")) if os.path.isdir(d)),
                                 key=lambda d: (not os.path.basename(d).isdigit(), int(os.path.basename(d)) if os.path.basename(d).isdigit() else 0))
        for sub in subs:
            s = summarise_folder(sub, args.skip)
            if s:
                report[os.path.relpath(sub)] = s
    phases = ("get_data", "data2dev", "zero_grad", "forward", "backprop", "sync", "optimizer_step")
    print(f"{'folder/experiment':58s} {'wrk':>3s} {'img/s':>10s} {'batch ms':>9s}  " + " ".join(f"{p[:8]:>8s}" for p in phases))
    for folder, exps in report.items():
        for name, r in exps.items():
            print(f"{(folder[-40:] + '/' + name)[:58]:58s} {r['workers']:3d} {r['images_per_sec']:10.1f} {r['batch_ms']:9.3f}  "
                  + " ".join(f"{r['phases_ms'].get(p, 0.0):8.3f}" for p in phases))
    if args.json:
        with open(args.json, "w") as f:
            json.dump(report, f, indent=1)


if __name__ == "__main__":
    main()
