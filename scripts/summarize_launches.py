"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel launches, total time, share."""
import csv
import io
import sys
from collections import defaultdict


def load(path):
    lines = [l for l in open(path) if l.startswith('"')]
    return list(csv.DictReader(io.StringIO("".join(lines))))


def main():
    path = sys.argv[1]
    d = defaultdict(lambda: [0, 0.0])
    for r in load(path):
        k = r["Kernel Name"][:100]
        d[k][0] += 1
        d[k][1] += float(r["Metric Value"])
    tot = sum(v[1] for v in d.values()) or 1.0
    print("| kernel | launches | total ms | share |\n|---|---|---|---|")
    for k, v in sorted(d.items(), key=lambda kv: -kv[1][1]):
        print(f"| `{k}` | {v[0]} | {v[1] / 1e6:.3f} | {v[1] / tot * 100:.1f}% |")


if __name__ == "__main__":
    main()
