#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel count, total and share.

usage: python tools/summarise_launches.py gpurun_out/launches.csv [--skip-setup] > profiles/<name>.md
"""
import csv
import re
import sys
from collections import OrderedDict


def short(name):
    m = re.search(r"(k_\w+)", name)
    base = m.group(1) if m else name.split("(")[0][-60:]
    if "torch" in name or "at::native" in name:
        base = "torch:" + (re.search(r"(\w+_kernel\w*)", name).group(1) if re.search(r"(\w+_kernel\w*)", name) else "kernel")
    tag = ""
    for c in ("alt_bn128", "bls12_377", "bw6_761"):
        if c in name:
            tag = c
            break
    if "Fp2" in name:
        tag += " Fp2"
    return base + (f" <{tag}>" if tag else "")


def main():
    path = sys.argv[1]
    rows = []
    with open(path) as f:
        lines = [ln for ln in f if ln.startswith('"')]
    for r in csv.DictReader(lines):
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        ns = float(r["Metric Value"].replace(",", ""))
        if r["Metric Unit"] in ("us", "usecond"):
            ns *= 1e3
        elif r["Metric Unit"] in ("ms", "msecond"):
            ns *= 1e6
        rows.append((int(r["ID"]), short(r["Kernel Name"]), r["Grid Size"], r["Block Size"], ns))
    agg = OrderedDict()
    for _, k, g, b, ns in rows:
        a = agg.setdefault(k, [0, 0.0, g, b])
        a[0] += 1
        a[1] += ns
    total = sum(a[1] for a in agg.values())
    print(f"source: {path}  ({len(rows)} launches, {total / 1e6:.3f} ms summed; ncu per-launch times are cold-cache and serialised)")
    print()
    print("| kernel | launches | total ms | avg us | share | grid | block |")
    print("|---|---:|---:|---:|---:|---|---|")
    for k, (n, ns, g, b) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"| `{k}` | {n} | {ns / 1e6:.3f} | {ns / n / 1e3:.1f} | {100 * ns / total:.1f}% | {g} | {b} |")


if __name__ == "__main__":
    main()
