"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel name."""
import csv, collections, re, sys


def load(path):
    rows = []
    with open(path) as f:
        lines = [l for l in f if not l.startswith("==")]
    for row in csv.DictReader(lines):
        if row.get("Metric Name") == "gpu__time_duration.sum":
            v = float(row["Metric Value"].replace(",", "")); u = row["Metric Unit"]
            us = v / 1000 if u in ("nsecond", "ns") else (v if u in ("usecond", "us") else v * 1000)
            name = re.sub(r"\(.*", "", row["Kernel Name"]).replace("void ", "")
            rows.append((re.sub(r"<.*", "", name), us))
    return rows


def main():
    rows = load(sys.argv[1])
    agg = collections.OrderedDict()
    for n, u in rows:
        c = agg.setdefault(n, [0, 0.0]); c[0] += 1; c[1] += u
    print(f"{len(rows)} launches, {sum(u for _, u in rows) / 1e3:.2f} ms serialized")
    for n, (c, u) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{n[:60]:60s} {c:4d} {u:10.1f} us  {u / c:8.1f} us/launch")


if __name__ == "__main__":
    main()
