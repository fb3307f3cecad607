#!/bin/bash
# usage (under gpurun): bash tools/ncu_traffic.sh <tag>: full ncu capture of one k_render launch of the DEFAULT bench workload
# (the launch bench.py's roofline is quoted on); the DRAM bytes of that launch go to profiles/render_traffic.json by hand.
tag=$1
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-lens-grad"
$CMD > gpurun_out/${tag}_traffic_plain.log 2>&1 &&
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:k_render -s 3 -c 1 -f -o gpurun_out/${tag}_render64 $CMD > gpurun_out/${tag}_traffic_ncu.log 2>&1
echo "full rc=$?"
M=smsp__inst_executed.sum,smsp__thread_inst_executed.sum,sm__inst_executed_pipe_fma.sum,sm__inst_executed_pipe_fmaheavy.sum,sm__inst_executed_pipe_fmalite.sum,sm__inst_executed_pipe_alu.sum,sm__inst_executed_pipe_lsu.sum,l1tex__t_sectors.sum,lts__t_sectors.sum,l1tex__t_sectors_pipe_lsu_mem_local_op_ld.sum,l1tex__t_sectors_pipe_lsu_mem_local_op_st.sum,l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum,dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum
timeout 1500 ncu --metrics $M --clock-control none -k regex:k_render -s 3 -c 1 --csv --log-file gpurun_out/${tag}_render64_counters.csv $CMD > gpurun_out/${tag}_traffic_ncu2.log 2>&1
echo "counters rc=$?"; ncu -i gpurun_out/${tag}_render64.ncu-rep --page raw --csv > gpurun_out/${tag}_render64_raw.csv 2>/dev/null
tail -1 gpurun_out/${tag}_traffic_plain.log | cut -c1-160
