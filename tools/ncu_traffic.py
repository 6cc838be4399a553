#!/usr/bin/env python
"""Summarises an `ncu --set full` capture of the training step for bench.py's `roofline.traffic`:
dram__bytes_read.sum + dram__bytes_write.sum per launch, grouped the way the library brackets its kernels
(head_fwd = k_tc_fwd*, head_bwd = k_tc_bwd_dz + k_tc_bwd_dp + k_reduce_partials, kld, prepare).

    ncu -i gpurun_out/r2_step_full.ncu-rep --page raw --csv > /tmp/raw.csv
    python tools/ncu_traffic.py /tmp/raw.csv tf32 r2_step_full.ncu-rep > profiles/r2_dram_traffic.json
"""
import csv
import json
import sys

GROUPS = (("head_fwd", ("k_tc_fwd", "k_simt_fwd")), ("head_bwd", ("k_tc_bwd", "k_reduce_partials", "k_simt_bwd")),
          ("kld", ("k_kld",)), ("prepare", ("k_step_prepare",)))
UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    hdr, units = rows[0], rows[1]
    col = {h: i for i, h in enumerate(hdr)}
    per_kernel = {}
    for r in rows[2:]:
        name = r[col["Kernel Name"]]
        tot = 0.0
        for m in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
            tot += float(r[col[m]].replace(",", "")) * UNIT[units[col[m]]]
        dur = float(r[col["gpu__time_duration.sum"]].replace(",", "")) * {"ns": 1e-3, "us": 1.0, "ms": 1e3}[units[col["gpu__time_duration.sum"]]]
        k = per_kernel.setdefault(name.split("(")[0].split("<")[0].split("::")[-1], {"bytes": [], "us": []})
        k["bytes"].append(tot)
        k["us"].append(dur)
    kern = {k: {"launches": len(v["bytes"]), "dram_bytes_per_launch": sum(v["bytes"]) / len(v["bytes"]),
                "us_per_launch_under_ncu": sum(v["us"]) / len(v["us"])} for k, v in per_kernel.items()}
    groups = {}
    for g, prefixes in GROUPS:
        tot = sum(v["dram_bytes_per_launch"] for k, v in kern.items() if k.startswith(prefixes))
        if tot:
            groups[g] = tot
    json.dump({"precision": sys.argv[2], "capture": sys.argv[3], "groups": groups, "kernels": kern,
               "note": "dram__bytes_read.sum + dram__bytes_write.sum per launch (cold-cache, serialised replays); a group "
                       "is the sum over the kernels the library times together"}, sys.stdout, indent=1)
    print()


if __name__ == "__main__":
    main()
