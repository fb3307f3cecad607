#!/usr/bin/env python3
"""profiles/render_traffic.json from an ncu report of one k_render launch of the DEFAULT bench workload.
usage: ncu -i gpurun_out/<tag>_render64.ncu-rep --page raw --csv > raw.csv ; python tools/ncu_to_traffic.py raw.csv[,more.csv] <source note> [spp w h]
(more.csv: the long-format csv of the --metrics pass of tools/ncu_traffic.sh; later files override earlier ones)
bench.py reads the file for roofline.traffic (DRAM bytes per launch) and for the per-stage counters of roofline_stages."""
import csv, json, sys
m = {}
for fn in sys.argv[1].split(","):
    rows = [r for r in csv.reader(open(fn)) if r and not r[0].startswith("==")]
    hdr = rows[0]
    if "Metric Name" in hdr:              # long format (ncu --metrics ... --csv): one row per metric, last kernel wins
        i_n, i_v = hdr.index("Metric Name"), hdr.index("Metric Value")
        for r in rows[1:]:
            try:
                m[r[i_n]] = float(r[i_v].replace(",", ""))
            except (ValueError, IndexError):
                pass
    else:                                 # raw page: header, units, one row per captured kernel
        for k, v in zip(hdr, rows[-1]):
            try:
                m[k] = float(v.replace(",", ""))
            except ValueError:
                pass
g = lambda k, d=0.0: m.get(k, d)
want = ["smsp__inst_executed.sum", "smsp__thread_inst_executed.sum", "sm__inst_executed_pipe_fma.sum", "sm__inst_executed_pipe_fmaheavy.sum",
        "sm__inst_executed_pipe_fmalite.sum", "sm__inst_executed_pipe_alu.sum", "sm__inst_executed_pipe_lsu.sum", "l1tex__t_sectors.sum",
        "lts__t_sectors.sum", "l1tex__t_sectors_pipe_lsu_mem_local_op_ld.sum", "l1tex__t_sectors_pipe_lsu_mem_local_op_st.sum",
        "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "dram__bytes_read.sum", "dram__bytes_write.sum"]
counters = {k: g(k) for k in want if k in m}
if "sm__inst_executed_pipe_fma.sum" not in counters:
    counters["sm__inst_executed_pipe_fma.sum"] = g("sm__inst_executed_pipe_fmaheavy.sum") + g("sm__inst_executed_pipe_fmalite.sum")
issue = g("smsp__issue_active.avg.pct_of_peak_sustained_active", g("sm__inst_issued.avg.pct_of_peak_sustained_active"))
out = {"spp_per_step": int(sys.argv[3]) if len(sys.argv) > 3 else 64, "width": int(sys.argv[4]) if len(sys.argv) > 4 else 3840,
       "height": int(sys.argv[5]) if len(sys.argv) > 5 else 2160,
       "dram_bytes_per_launch": g("dram__bytes_read.sum") + g("dram__bytes_write.sum"),
       "dram_bytes_read": g("dram__bytes_read.sum"), "dram_bytes_write": g("dram__bytes_write.sum"),
       "kernel_ms_under_ncu": g("gpu__time_duration.sum") / 1e6, "source": sys.argv[2] if len(sys.argv) > 2 else "",
       "issue_slot_utilisation": issue / 100.0,
       "active_lanes_per_instruction": g("smsp__thread_inst_executed_per_inst_executed.ratio"),
       "l1_hit_rate": g("l1tex__t_sector_hit_rate.pct") / 100.0, "l2_hit_rate": g("lts__t_sector_hit_rate.pct") / 100.0,
       "warps_active_pct": g("sm__warps_active.avg.pct_of_peak_sustained_active"),
       "top_stalls_per_issue": {k.split("issue_stalled_")[1].split("_per_issue_active")[0]: round(v, 3) for k, v in sorted(
           ((k, v) for k, v in m.items() if k.startswith("smsp__average_warps_issue_stalled_") and k.endswith("_per_issue_active.ratio")),
           key=lambda kv: -kv[1])[:6]},
       "counters": counters}
json.dump(out, open("profiles/render_traffic.json", "w"), indent=1)
print(json.dumps(out, indent=1))
