#!/bin/bash
# usage (under gpurun): bash tools/ncu_traffic.sh <tag>: full ncu capture of one k_render launch of the DEFAULT bench workload
# (the launch bench.py's roofline is quoted on); the DRAM bytes of that launch go to profiles/render_traffic.json by hand.
tag=$1
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-lens-grad"
$CMD > gpurun_out/${tag}_traffic_plain.log 2>&1 &&
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:k_render -s 3 -c 1 -f -o gpurun_out/${tag}_render64 $CMD > gpurun_out/${tag}_traffic_ncu.log 2>&1
echo "rc=$?"; tail -1 gpurun_out/${tag}_traffic_plain.log | cut -c1-160
