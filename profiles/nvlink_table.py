#!/usr/bin/env python3
"""Join the two single-pass ncu captures of rank 0 (durations; NVLink rx/tx bytes) of
profiles/nvlink_worker.py and print, per launch, the link bytes against the algorithmic bytes.

    python profiles/nvlink_table.py profiles/ncu/r2_nvlink_n2 2

Algorithmic bytes per direction per GPU for a buffer of S bytes on N ranks: two-shot 2 (N-1)/N S,
one-shot (N-1) S, nvls (N-1)/N S + S  (the peers' switch reductions pull (N-1)/N S of this GPU's data
and its own reduction sends S/N more; the multicast stores deliver S).  The counters include the
NVLink protocol overhead (headers, read requests), at 32-byte granularity.
"""
import csv
import json
import sys

SEQ = [("twoshot", 64 << 20), ("oneshot", 1 << 20), ("twoshot+pack+sgd", 16 << 20), ("nvls u8", 64 << 20), ("nvls u4", 4 << 20),
       ("nvls u8+pack+sgd", 16 << 20)]


def read(path):
    out = {}
    with open(path) as f:
        rows = [r for r in csv.reader(f) if len(r) > 10]
    for r in rows[1:]:
        out.setdefault(int(r[0]), {"kernel": r[4].split("(")[0].replace("void <unnamed>::", "")})[r[12]] = float(r[14])
    return out


def main():
    prefix, n = sys.argv[1], int(sys.argv[2])
    t, b = read(prefix + "_durations.csv"), read(prefix + "_bytes.csv")
    rows = []
    for i in sorted(t):
        name, size = SEQ[i]
        alg = {"two": 2 * (n - 1) / n * size, "one": (n - 1) * size, "nvl": (n - 1) / n * size + size}[name[:3]]
        us = t[i]["gpu__time_duration.sum"] / 1000.0
        rx, tx = b[i]["nvlrx__bytes.sum"], b[i]["nvltx__bytes.sum"]
        rows.append({"launch": name, "kernel": t[i]["kernel"], "MiB": size >> 20, "us_under_ncu": round(us, 1),
                     "nvl_rx_MB": round(rx / 1e6, 2), "nvl_tx_MB": round(tx / 1e6, 2), "algorithmic_MB": round(alg / 1e6, 2),
                     "rx_over_algorithmic": round(rx / alg, 3), "rx_GBs": round(rx / us / 1e3, 1), "tx_GBs": round(tx / us / 1e3, 1)})
    print(json.dumps({"n_gpus": n, "rows": rows}, indent=1))


if __name__ == "__main__":
    main()
