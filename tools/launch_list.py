"""Summarise an ncu launch list (`ncu --metrics gpu__time_duration.sum --clock-control none
--csv --log-file X.csv python bench.py ...`) per kernel: launches, total time, share.

    python tools/launch_list.py gpurun_out/r02_launches.csv "<command>" > profiles/r02_launches.txt
"""
import csv
import sys
from collections import OrderedDict


def main():
    path, cmd = sys.argv[1], (sys.argv[2] if len(sys.argv) > 2 else "")
    rows = [r for r in csv.reader(l for l in open(path) if not l.startswith("=="))]
    hdr = rows[0]
    ik, im, iv, iu = hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    agg = OrderedDict()
    for r in rows[1:]:
        if len(r) <= iv or r[im] != "gpu__time_duration.sum":
            continue
        v = float(r[iv].replace(",", ""))
        v *= {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(r[iu], 1e-6)
        n, t = agg.get(r[ik], (0, 0.0))
        agg[r[ik]] = (n + 1, t + v)
    tot = sum(t for _, t in agg.values()) or 1.0
    print(f"# ncu --metrics gpu__time_duration.sum --clock-control none: {cmd}")
    print("# per-launch times are cold-cache and serialised under ncu: read the SHARE, not the absolute")
    for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{k[:80]:<82} launches {n:4d}  total {t:9.3f} ms  share {100 * t / tot:5.1f} %")


if __name__ == "__main__":
    main()
