"""Summarise an `ncu --metrics gpu__time_duration.sum[,...] --csv` launch list: one wavefront pass, per kernel."""
import collections
import csv
import re
import sys


def load(path):
    rows = list(csv.reader(open(path)))
    hdr = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
    cols = rows[hdr]
    ki, mi, vi, ui, ii = cols.index("Kernel Name"), cols.index("Metric Name"), cols.index("Metric Value"), cols.index("Metric Unit"), cols.index("ID")
    launches = collections.OrderedDict()
    for r in rows[hdr + 1:]:
        if len(r) <= vi:
            continue
        name = re.sub(r"^.*::", "", re.sub(r"\(.*", "", r[ki]))
        v = float(r[vi].replace(",", ""))
        if r[mi] == "gpu__time_duration.sum":
            v = v / 1e3 if r[ui] in ("ns", "nsecond") else (v * 1e3 if r[ui] in ("ms", "msecond") else v)
        launches.setdefault(r[ii], {"name": name})[r[mi]] = v
    return list(launches.values())


if __name__ == "__main__":
    seq = load(sys.argv[1])
    which = int(sys.argv[2]) if len(sys.argv) > 2 else -2
    starts = [i for i, l in enumerate(seq) if l["name"].startswith("raygen") or l["name"].startswith("primary_packet")]
    s0 = starts[which]
    s1 = starts[which + 1] if which + 1 < len(starts) and which != -1 else len(seq)
    tot = collections.Counter()
    print(f"{len(seq)} launches captured; pass starting at launch {s0}:")
    for l in seq[s0:s1]:
        extra = "  ".join(f"{k.split('.')[0].split('__')[-1][:28]}={v:.2f}" for k, v in l.items() if k not in ("name", "gpu__time_duration.sum"))
        print(f"  {l['name']:28s} {l['gpu__time_duration.sum']:9.1f} us  {extra}")
        tot[l["name"]] += l["gpu__time_duration.sum"]
    total = sum(tot.values())
    for k, v in tot.most_common():
        print(f"  share {k:28s} {v:9.1f} us  {100 * v / total:5.1f} %")
    print(f"  pass total {total:.1f} us")
