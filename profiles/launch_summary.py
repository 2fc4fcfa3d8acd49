"""Share-of-time table from an ncu launch list (`ncu --metrics gpu__time_duration.sum --clock-control none --csv`).
    python profiles/launch_summary.py <launches.csv> ["header line"] > summary.txt
Times under ncu are cold-cache and serialised: compare SHARES, not absolutes."""
import csv
import re
import sys
from collections import defaultdict


def short(name):
    name = re.sub(r"^void ", "", name)
    name = re.sub(r"\(.*$", "", name)
    name = name.replace("p2p::", "").replace("(anonymous namespace)::", "").replace("fast::", "")
    return name[:90]


def main(path, header=None):
    rows = [r for r in csv.reader(l for l in open(path, errors="replace") if l.startswith('"'))]
    hdr = rows[0]
    ik, iv, iu = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    tot, cnt = defaultdict(float), defaultdict(int)
    for r in rows[1:]:
        v = float(r[iv].replace(",", ""))
        us = v / 1e3 if r[iu] in ("ns", "nsecond") else (v if r[iu] in ("us", "usecond") else v * 1e3)
        k = short(r[ik])
        tot[k] += us
        cnt[k] += 1
    total = sum(tot.values())
    if header:
        print(header)
    print("total %.1f us over %d launches\n" % (total, sum(cnt.values())))
    for k in sorted(tot, key=tot.get, reverse=True)[:40]:
        print("%6.2f%% %10.1f us %6d launches  %s" % (100 * tot[k] / total, tot[k], cnt[k], k))


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else None)
