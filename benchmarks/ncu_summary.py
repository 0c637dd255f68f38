"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: time and count per kernel."""
import collections
import csv
import sys


def main(path, top=30):
    with open(path) as f:
        lines = [ln for ln in f if not ln.startswith("==")]
    tot, cnt = collections.defaultdict(float), collections.Counter()
    for row in csv.DictReader(lines):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(row["Metric Value"].replace(",", ""))
        unit = row["Metric Unit"]
        v = v / 1e3 if unit == "ns" else (v * 1e3 if unit == "ms" else v)
        key = row["Kernel Name"].split("(")[0].replace("void ", "").replace("<unnamed>::", "")[:90]
        tot[key] += v
        cnt[key] += 1
    print(f"{sum(cnt.values())} launches, {sum(tot.values()) / 1e3:.2f} ms total")
    for k, v in sorted(tot.items(), key=lambda kv: -kv[1])[:top]:
        print(f"{v:9.1f} us {cnt[k]:5d}  {k}")


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 30)
