#!/usr/bin/env python
"""Summarise `ncu --set full` captures (one kernel launch each) into a small JSON for profiles/.
usage: python tools/ncu_summary.py out.json rep1.ncu-rep [rep2 ...]"""
import csv, json, subprocess, sys

KEEP = ["gpu__time_duration.sum", "gpc__cycles_elapsed.avg.per_second", "dram__cycles_elapsed.avg.per_second",
        "smsp__cycles_elapsed.avg.per_second", "dram__bytes_read.sum", "dram__bytes_write.sum", "sm__cycles_elapsed.max",
        "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "launch__grid_size", "launch__block_size",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "lts__t_sectors_op_red.sum", "lts__t_sectors_op_atom.sum",
        "lts__t_bytes.sum", "l1tex__t_bytes.sum"]
UNIT = {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1.0}


def one(rep):
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(txt.splitlines()))
    h, u = rows[0], rows[1]
    out = []
    for v in rows[2:]:
        d = {"kernel": v[h.index("Kernel Name")]}
        for i, k in enumerate(h):
            if k in KEEP:
                try:
                    x = float(v[i])
                except ValueError:
                    continue
                if u[i] in UNIT:
                    x *= UNIT[u[i]]; d[k + " [byte]"] = x
                else:
                    d[k + (f" [{u[i]}]" if u[i] else "")] = x
        rd, wr = d.get("dram__bytes_read.sum [byte]"), d.get("dram__bytes_write.sum [byte]")
        if rd is not None and wr is not None:
            d["traffic_bytes"] = rd + wr
        out.append(d)
    return out


if __name__ == "__main__":
    res = {}
    for rep in sys.argv[2:]:
        res[rep.split("/")[-1]] = one(rep)
    json.dump(res, open(sys.argv[1], "w"), indent=1)
    for k, v in res.items():
        for d in v:
            print(k, d["kernel"][:60], d.get("gpu__time_duration.sum [ms]", d.get("gpu__time_duration.sum [us]")), d.get("traffic_bytes"))
