"""Aggregates an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel name (share of the summed,
serialised, cold-cache device time). Usage: python profiles/summarize_launches.py gpurun_out/launches.csv"""
import collections
import csv
import re
import sys


def main(path):
    with open(path) as f:
        lines = [l for l in f if not l.startswith("==")]
    agg = collections.defaultdict(lambda: [0, 0.0])
    tot = 0.0
    for row in csv.DictReader(lines):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(row["Metric Value"].replace(",", ""))
        unit = row["Metric Unit"]
        v = v / 1000 if unit == "ns" else (v * 1000 if unit == "ms" else v)
        name = re.sub(r"\(.*", "", row["Kernel Name"])
        name = name.replace("fqlpop::<unnamed>::", "").replace("void ", "")[:70]
        agg[name][0] += 1
        agg[name][1] += v
        tot += v
    print(f"{'kernel':70s} {'launches':>8s} {'total_us':>10s} {'avg_us':>8s} {'share':>6s}")
    for k, (n, t) in sorted(agg.items(), key=lambda x: -x[1][1]):
        print(f"{k:70s} {n:8d} {t:10.1f} {t / n:8.2f} {t / tot:6.3f}")
    print(f"{'TOTAL':70s} {'':8s} {tot:10.1f}")


if __name__ == "__main__":
    main(sys.argv[1])
