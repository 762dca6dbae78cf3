"""Reduce an ncu launch list (`ncu --metrics gpu__time_duration.sum --clock-control none --csv
--log-file X.csv <command>`) to per-kernel totals and shares.
Usage: python profiles/summarize_launches.py X.csv out.json "<the command that was profiled>" """
import collections
import csv
import json
import sys


def main():
    src, dst, cmd = sys.argv[1], sys.argv[2], sys.argv[3] if len(sys.argv) > 3 else ""
    rows = []
    with open(src, newline="") as f:
        lines = [ln for ln in f if not ln.startswith("==")]
    rd = csv.DictReader(lines)
    for r in rd:
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(r["Metric Value"].replace(",", ""))
        unit = r.get("Metric Unit", "ns")
        us = v / 1e3 if unit in ("ns", "nsecond") else v * 1e3 if unit in ("ms", "msecond") else v
        rows.append((r["Kernel Name"], us))
    tot = collections.OrderedDict()
    for k, us in rows:
        e = tot.setdefault(k[:110], [0, 0.0])
        e[0] += 1
        e[1] += us
    total = sum(e[1] for e in tot.values()) or 1.0
    out = {"command": cmd,
           "note": "cold-cache, serialised per-launch times: the shares are what must agree with the bench line",
           "kernels": [{"kernel": k, "launches": n, "total_us": round(us, 1), "us_per_launch": round(us / n, 1),
                        "share_of_listed_kernels": round(us / total, 4)} for k, (n, us) in tot.items()]}
    json.dump(out, open(dst, "w"), indent=1)
    print(json.dumps(out)[:600])


if __name__ == "__main__":
    main()
