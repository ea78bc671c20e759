#!/usr/bin/env python
"""Summarises an `ncu --set full` report of the accumulation kernels into profiles/:

    python tools/ncu_summary.py gpurun_out/prof_fwd_bwd.ncu-rep profiles/r01_ncu_full_fwd_bwd_v3.json

Writes the selected raw metrics per kernel launch and updates profiles/ncu_traffic.json
({kernel base name: dram bytes read+written per launch}), which bench.py reports as roofline.traffic."""
import csv
import io
import json
import os
import re
import subprocess
import sys

KEEP = ("Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput",
        "sm__throughput.avg.pct", "smsp__inst_executed.sum", "launch__registers_per_thread", "launch__grid_size",
        "launch__block_size", "launch__occupancy_limit", "launch__waves_per_multiprocessor",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__average_warps_issue_stalled", "sm__inst_executed_pipe_fma.avg", "sm__inst_executed_pipe_xu.avg",
        "sm__inst_executed_pipe_alu.avg", "sm__inst_executed_pipe_lsu.avg", "sm__pipe_fma_cycles_active.avg",
        "sm__pipe_fmaheavy_cycles_active.avg", "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct",
        "lts__t_bytes.sum ", "l1tex__data_bank_conflicts_pipe_lsu.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "sm__cycles_elapsed.avg ", "smsp__cycles_active.avg ")


def main():
    rep, out = sys.argv[1], sys.argv[2]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    keep = [i for i, h in enumerate(hdr) if any((h + " ").startswith(k) or h.startswith(k) for k in KEEP)]
    res, traffic = [], {}
    for r in rows[2:]:
        d = {}
        for i in keep:
            d[hdr[i] + (f" [{units[i]}]" if units[i] else "")] = r[i]
        res.append(d)
        name = re.sub(r"^void ", "", r[hdr.index("Kernel Name")])
        base = re.sub(r"[<(].*", "", name)

        def val(metric):
            i = hdr.index(metric)
            v = float(r[i].replace(",", ""))
            u = units[i].lower()
            return v * {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}.get(u, 1)
        traffic[base] = val("dram__bytes_read.sum") + val("dram__bytes_write.sum")
    json.dump(res, open(out, "w"), indent=1)
    tpath = os.path.join(os.path.dirname(out) or ".", "ncu_traffic.json")
    old = json.load(open(tpath)) if os.path.exists(tpath) else {}
    old.update(traffic)
    old["_source"] = os.path.basename(out)
    json.dump(old, open(tpath, "w"), indent=1)
    for d in res:
        print({k: v for k, v in d.items() if k.startswith(("Kernel", "gpu__time", "dram__bytes", "smsp__inst_executed.sum",
                                                            "smsp__issue_active", "launch__registers"))})


if __name__ == "__main__":
    main()
