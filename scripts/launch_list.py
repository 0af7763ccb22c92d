"""Print an ncu launch list (--metrics gpu__time_duration.sum --csv) kernel by kernel:
    python scripts/launch_list.py FILE.csv [launches-per-batch-group | 0 = totals by kernel]"""
import collections
import csv
import re
import sys


def read(path):
    with open(path) as f:
        lines = [l for l in f if not l.startswith("==")]
    seq = []
    for row in csv.DictReader(lines):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        name = re.sub(r"\(.*", "", row["Kernel Name"]).replace("b200r::", "")
        v = float(row["Metric Value"].replace(",", ""))
        v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(row["Metric Unit"], 1.0)
        seq.append((name, v))
    return seq


if __name__ == "__main__":
    seq = read(sys.argv[1])
    last = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    if last:
        for n, v in seq[-last:]:
            print("%-44s %9.1f us" % (n[:44], v))
        print("%-44s %9.1f us" % ("sum", sum(v for _, v in seq[-last:])))
    else:
        tot, cnt = collections.OrderedDict(), collections.Counter()
        for n, v in seq:
            tot[n] = tot.get(n, 0.0) + v
            cnt[n] += 1
        for n, v in tot.items():
            print("%-44s x%-4d mean %9.1f us  total %10.1f us" % (n[:44], cnt[n], v / cnt[n], v))
