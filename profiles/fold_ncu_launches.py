"""Fold an ncu launch list (--csv, metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum) into
per-kernel totals for the LAST optimizer step in the capture (from the launch after the previous adamw_flat_kernel to
the last adamw_flat_kernel, inclusive).

    python profiles/fold_ncu_launches.py profiles/r01c_ncu_launches_bench_value_loop.csv > profiles/r01c_ncu_launch_list_step.json
"""
import csv
import json
import re
import sys


def short(name):
    name = re.sub(r"^void ", "", name)
    name = re.sub(r"^penroz::", "", name)
    m = re.match(r"([A-Za-z0-9_:]+)(<[^(]*>)?\(", name)
    if not m:
        return name[:60]
    base, targs = m.group(1), m.group(2) or ""
    if base.startswith("gemm_bf16_tc") or base.startswith("gemv"):
        return base.replace("_kernel", "") + targs.replace("penroz::", "")
    return base


def main(path):
    launches = {}
    order = []
    with open(path, newline="") as f:
        rows = [r for r in csv.reader(f) if len(r) > 14 and r[0].isdigit()]
    for r in rows:
        i = int(r[0])
        if i not in launches:
            launches[i] = {"name": r[4]}
            order.append(i)
        launches[i][r[12]] = float(r[14].replace(",", "")) * {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "byte": 1.0, "Kbyte": 1e3,
                                                               "Mbyte": 1e6, "Gbyte": 1e9}.get(r[13], 1.0)
    adam = [i for i in order if "adamw_flat" in launches[i]["name"]]
    if len(adam) < 2:
        raise SystemExit("need at least two optimizer steps in the capture")
    # with bucket-wise AdamW several adamw launches end a step: a step ends at the last adamw of a run of them
    ends = [a for k, a in enumerate(adam) if k + 1 == len(adam) or adam[k + 1] != order[order.index(a) + 1]]
    lo, hi = order.index(ends[-2]) + 1, order.index(ends[-1])
    fam = {}
    for i in order[lo:hi + 1]:
        L = launches[i]
        k = short(L["name"])
        d = fam.setdefault(k, {"kernel": k, "launches": 0, "ms": 0.0, "dram_read_MB": 0.0, "dram_write_MB": 0.0})
        d["launches"] += 1
        d["ms"] += L.get("gpu__time_duration.sum", 0.0)
        d["dram_read_MB"] += L.get("dram__bytes_read.sum", 0.0) / 1e6
        d["dram_write_MB"] += L.get("dram__bytes_write.sum", 0.0) / 1e6
    total = sum(d["ms"] for d in fam.values())
    fams = sorted(fam.values(), key=lambda d: -d["ms"])
    for d in fams:
        d["share"] = round(d["ms"] / total, 4)
        for k in ("ms", "dram_read_MB", "dram_write_MB"):
            d[k] = round(d[k], 4 if k == "ms" else 1)
    print(json.dumps({"source": "ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum "
                                "--clock-control none on `PENROZ_BENCH_PARTS=value python bench.py --steps 2 --warmup 3`; "
                                "last optimizer step", "step_ms_serialised": round(total, 3),
                      "launches": hi - lo + 1, "families": fams}, indent=1))


if __name__ == "__main__":
    main(sys.argv[1])
