#!/usr/bin/env python3
"""Selected raw metrics of one kernel launch from an ncu report, as text and as JSON.
usage: tools/ncu_summary.py prof.ncu-rep "header line" out.txt [out.json]"""
import csv, json, subprocess, sys
rep, header, out_txt = sys.argv[1:4]
out_json = sys.argv[4] if len(sys.argv) > 4 else None
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
names, units, vals = rows[0], rows[1], rows[2]
d = {n: (u, v) for n, u, v in zip(names, units, vals)}
want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "launch__registers_per_thread", "launch__grid_size",
        "launch__block_size", "launch__shared_mem_per_block_dynamic", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "smsp__inst_executed_op_shared_atom.sum",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "l1tex__t_sectors_pipe_lsu_mem_local_op_ld.sum", "l1tex__t_sectors_pipe_lsu_mem_local_op_st.sum",
        "sm__cycles_active.avg", "sm__cycles_active.max", "sm__cycles_active.min", "lts__t_sector_hit_rate.pct"]
want += sorted(k for k in d if "issue_stalled" in k and k.endswith("per_issue_active.ratio") and "not_issued" not in k)
with open(out_txt, "w") as f:
    f.write(header + "\n")
    for k in want:
        if k in d:
            f.write("%s [%s] = %s\n" % (k, d[k][0], d[k][1]))
if out_json:
    def num(k):
        u, v = d[k]
        x = float(v)
        mult = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}.get(u, 1.0)
        return x * mult
    rd, wr = num("dram__bytes_read.sum"), num("dram__bytes_write.sum")
    u, v = d["gpu__time_duration.sum"]
    ms = float(v) * {"ms": 1.0, "us": 1e-3, "s": 1e3, "ns": 1e-6}.get(u, 1.0)
    def plain(k):
        return float(d[k][1]) if k in d else None
    json.dump({"source": rep.split("/")[-1], "kernel": d.get("Kernel Name", ("", ""))[1], "dram_bytes_read": rd,
               "dram_bytes_write": wr, "traffic_bytes_per_launch": rd + wr, "time_ms_under_ncu": ms,
               "warp_instructions_per_launch": plain("smsp__inst_executed.sum"),
               "shared_wavefronts_per_launch": plain("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum"),
               "shared_bank_conflict_wavefronts_per_launch": plain("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum"),
               "sm_cycles_active_avg": plain("sm__cycles_active.avg")}, open(out_json, "w"), indent=1)
    open(out_json, "a").write("\n")
