"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel (shares of the step)."""
import collections
import csv
import re
import sys


def main(path, note=""):
    rows = list(csv.reader(l for l in open(path) if l.startswith('"')))
    hdr = rows[0]
    ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    agg = collections.defaultdict(lambda: [0, 0.0])
    for r in rows[1:]:
        name = re.sub(r'\(.*', '', r[ki])[:90]
        v = float(r[vi].replace(',', ''))
        v = v / 1000.0 if r[ui] == 'ns' else v * 1000.0 if r[ui] == 'ms' else v
        agg[name][0] += 1
        agg[name][1] += v
    tot = sum(v[1] for v in agg.values())
    if note:
        print("# " + note)
    print("# per-launch times are cold-cache and serialised under ncu: compare SHARES, not absolutes")
    print(f"# total {tot / 1000:.1f} ms over {sum(v[0] for v in agg.values())} launches")
    print(f"{'sum_us':>12} {'share':>6} {'n':>6} {'avg_us':>9}  kernel")
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:30]:
        print(f"{v[1]:12.1f} {100 * v[1] / tot:5.1f}% {v[0]:6d} {v[1] / v[0]:9.1f}  {k}")


if __name__ == "__main__":
    main(sys.argv[1], " ".join(sys.argv[2:]))
