"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel:  python profiles/summarize_launches.py x.csv [skip_first_n]"""
import collections
import csv
import sys


def main(path, skip=0):
    rows = list(csv.reader(open(path)))
    for i, r in enumerate(rows):
        if "Kernel Name" in r:
            h, start = r, i + 1
            break
    ix = {k: j for j, k in enumerate(h)}
    agg = collections.OrderedDict()
    n = 0
    for r in rows[start:]:
        if len(r) < len(h) or r[ix["Metric Name"]] != "gpu__time_duration.sum":
            continue
        n += 1
        if n <= skip:
            continue
        v = float(r[ix["Metric Value"]].replace(",", ""))
        u = r[ix["Metric Unit"]]
        v = v / 1000 if u in ("ns", "nsecond") else (v * 1000 if u in ("ms", "msecond") else v)
        key = (r[ix["Kernel Name"]][:70], r[ix["Grid Size"]], r[ix["Block Size"]])
        a = agg.setdefault(key, [0, 0.0])
        a[0] += 1
        a[1] += v
    tot = sum(a[1] for a in agg.values())
    print(f"{n - skip} launches, {tot:.1f} us total (cold-cache, serialised: compare shares)\n")
    print("| us total | launches | us each | share | kernel | grid | block |\n|---|---|---|---|---|---|---|")
    for (k, g, b), (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"| {t:.1f} | {c} | {t / c:.2f} | {100 * t / tot:.1f}% | `{k}` | {g} | {b} |")


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 0)
