#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list into a per-kernel share table (markdown)."""
import collections
import csv
import re
import sys


def main(path, title):
    with open(path) as f:
        lines = [l for l in f if not l.startswith("==")]
    tot, cnt = collections.defaultdict(float), collections.Counter()
    for row in csv.DictReader(lines):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        kn = row["Kernel Name"]
        m = re.search(r"(\w+_kernel|col_reduce_stage\d)", kn)
        name = m.group(1) if m else kn[:50]
        t = re.search(re.escape(name) + r"<([^>]*)>", kn)
        if t and ("gemm" in name):
            name += "<" + t.group(1).replace("(int)", "") + ">"
        v = float(row["Metric Value"].replace(",", ""))
        v *= {"ns": 1.0, "us": 1e3, "ms": 1e6, "nsecond": 1.0, "usecond": 1e3, "msecond": 1e6}.get(row["Metric Unit"], 1.0)
        tot[name] += v
        cnt[name] += 1
    T = sum(tot.values())
    print(f"# {title}\n")
    print(f"Source: `{path}` (ncu, `--metrics gpu__time_duration.sum --clock-control none`; per-launch times are cold-cache and")
    print(f"serialised: compare SHARES, not absolutes). {sum(cnt.values())} launches, {T / 1e6:.2f} ms summed.\n")
    print("| kernel | launches | total ms | share | avg us |\n|---|---:|---:|---:|---:|")
    for k, v in sorted(tot.items(), key=lambda kv: -kv[1]):
        print(f"| `{k}` | {cnt[k]} | {v / 1e6:.3f} | {100 * v / T:.1f}% | {v / cnt[k] / 1e3:.1f} |")


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else "kernel launch list")
