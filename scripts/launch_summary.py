"""Aggregate an ncu `--metrics gpu__time_duration.sum --csv` launch list per (kernel, grid)."""
import collections
import csv
import re
import sys


def main(path, steps=2):
    with open(path) as f:
        lines = [l for l in f if not l.startswith("==")]
    agg, tot = collections.OrderedDict(), 0.0
    for row in csv.DictReader(lines):
        name = re.sub(r"\(.*", "", row["Kernel Name"]).replace("void ", "")
        v = float(row["Metric Value"].replace(",", ""))
        v *= {"ns": 1, "us": 1e3, "ms": 1e6}.get(row["Metric Unit"], 1)
        a = agg.setdefault((name, row["Grid Size"], row["Block Size"]), [0, 0.0])
        a[0] += 1
        a[1] += v
        tot += v
    print("| kernel | grid | block | launches | avg us | share |\n|---|---|---|---|---|---|")
    for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"| {k[0]} | {k[1]} | {k[2]} | {n} | {t / n / 1e3:.2f} | {100 * t / tot:.1f}% |")
    print(f"\nTotal {tot / 1e3:.0f} us over {steps} steps = {tot / steps / 1e3:.0f} us/step under ncu.")


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 2)
