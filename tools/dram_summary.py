#!/usr/bin/env python
"""Summarise an ncu `--metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --csv`
launch list of ONE training step: per kernel family launches, time, DRAM bytes and achieved DRAM GB/s
(under ncu: cold cache, serialised -- shares and traffic matter, not absolute times).

    python tools/dram_summary.py gpurun_out/dram_step.csv [--json out.json]
"""
import collections
import csv
import json
import re
import sys


def load(path):
    rows = list(csv.reader(open(path)))
    hi = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
    hdr = {h: i for i, h in enumerate(rows[hi])}
    per = collections.OrderedDict()
    for r in rows[hi + 1:]:
        if len(r) < len(hdr):
            continue
        k = per.setdefault(int(r[hdr["ID"]]), {"name": r[hdr["Kernel Name"]], "grid": r[hdr["Grid Size"]]})
        v, unit = float(r[hdr["Metric Value"]].replace(",", "")), r[hdr["Metric Unit"]]
        m = r[hdr["Metric Name"]]
        if m.startswith("gpu__time"):
            v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6, "nsecond": 1e-3, "usecond": 1.0, "msecond": 1e3}.get(unit, 1.0)
        else:
            v *= {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1.0)
        k[m] = v
    return list(per.values())


def family(name):
    name = re.sub(r"^void\s+", "", name)
    name = re.sub(r"\(anonymous namespace\)::|<unnamed>::", "", name)
    m = re.match(r"([A-Za-z0-9_:]+)", name)
    return m.group(1) if m else name


def main():
    path = sys.argv[1]
    ks = load(path)
    fam = collections.OrderedDict()
    for k in ks:
        f = fam.setdefault(family(k["name"]), {"launches": 0, "us": 0.0, "rd": 0.0, "wr": 0.0})
        f["launches"] += 1
        f["us"] += k.get("gpu__time_duration.sum", 0.0)
        f["rd"] += k.get("dram__bytes_read.sum", 0.0)
        f["wr"] += k.get("dram__bytes_write.sum", 0.0)
    tot_us = sum(f["us"] for f in fam.values())
    print(f"{len(ks)} launches, {tot_us / 1e3:.3f} ms summed kernel time (ncu, serialised)")
    print(f"{'kernel':44s} {'n':>4s} {'ms':>8s} {'share':>6s} {'read MB':>9s} {'write MB':>9s} {'GB/s':>7s}")
    for name, f in sorted(fam.items(), key=lambda kv: -kv[1]["us"]):
        gbs = (f["rd"] + f["wr"]) / max(f["us"], 1e-9) * 1e-3
        print(f"{name[:44]:44s} {f['launches']:4d} {f['us'] / 1e3:8.3f} {100 * f['us'] / tot_us:5.1f}% "
              f"{f['rd'] / 1e6:9.1f} {f['wr'] / 1e6:9.1f} {gbs:7.0f}")
    if "--json" in sys.argv:
        out = sys.argv[sys.argv.index("--json") + 1]
        json.dump({"source": path, "launches": len(ks), "kernel_ms": tot_us / 1e3,
                   "families": {n: f for n, f in fam.items()}}, open(out, "w"), indent=1)


if __name__ == "__main__":
    main()
