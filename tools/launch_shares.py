"""Per-kernel totals of an `ncu --metrics gpu__time_duration.sum --csv` launch list."""
import collections
import csv
import re
import sys


def main(path):
    rows = list(csv.reader(open(path)))
    hi = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
    hdr = rows[hi]
    kn, mv = hdr.index("Kernel Name"), hdr.index("Metric Value")
    agg = collections.defaultdict(lambda: [0, 0.0])
    for r in rows[hi + 1:]:
        if len(r) <= mv:
            continue
        try:
            v = float(r[mv].replace(",", ""))
        except ValueError:
            continue
        name = re.sub(r"\(.*", "", r[kn])
        agg[name][0] += 1
        agg[name][1] += v
    tot = sum(v for _, v in agg.values())
    print("%d launches, %.1f us" % (sum(c for c, _ in agg.values()), tot / 1000))
    for n, (c, v) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print("%-64s %4d %9.1f us %5.1f%%  %8.1f us/launch" % (n[:64], c, v / 1000, 100 * v / tot, v / c / 1000))


if __name__ == "__main__":
    main(sys.argv[1])
