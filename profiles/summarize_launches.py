"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel.

    python profiles/summarize_launches.py gpurun_out/launches.csv > profiles/<name>.txt

Per-launch times under ncu are cold-cache and serialised: compare SHARES of the step, not absolutes."""
import collections
import csv
import re
import sys


def main(path, top=40):
    with open(path) as fh:
        lines = [l for l in fh if not l.startswith("==")]
    total, count = collections.defaultdict(float), collections.Counter()
    for row in csv.DictReader(lines):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        value = float(row["Metric Value"].replace(",", ""))
        value *= {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(row["Metric Unit"], 1.0)
        name = re.sub(r"^void ", "", row["Kernel Name"])
        name = re.sub(r"[<(].*", "", name)[:80]
        total[name] += value
        count[name] += 1
    grand = sum(total.values())
    print(f"# {path}: {sum(count.values())} launches, {grand / 1e3:.2f} ms of kernel time")
    print(f"{'share':>7} {'total_us':>11} {'launches':>8} {'avg_us':>9}  kernel")
    for name, t in sorted(total.items(), key=lambda kv: -kv[1])[:top]:
        print(f"{t / grand * 100:6.2f}% {t:11.1f} {count[name]:8d} {t / count[name]:9.1f}  {name}")


if __name__ == "__main__":
    main(sys.argv[1])
