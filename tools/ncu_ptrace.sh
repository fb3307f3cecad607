#!/bin/bash
# usage (under gpurun): bash tools/ncu_ptrace.sh <tag>: full ncu capture of one k_ptrace launch of bench.py's particle-tracer block
# (headline scene and lens, 1024x576 film, 4 spp) + the launch's counters; summaries go to profiles/ by hand.
tag=$1
CMD="python tools/ptrace_once.py"
$CMD > gpurun_out/${tag}_ptrace_plain.log 2>&1; tail -1 gpurun_out/${tag}_ptrace_plain.log | cut -c1-200
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:k_ptrace -s 4 -c 1 -f -o gpurun_out/${tag}_ptrace $CMD > gpurun_out/${tag}_ptrace_ncu.log 2>&1
echo "full rc=$?"
ncu -i gpurun_out/${tag}_ptrace.ncu-rep --page raw --csv > gpurun_out/${tag}_ptrace_raw.csv 2>/dev/null
ncu -i gpurun_out/${tag}_ptrace.ncu-rep --page source --print-source cuda,sass --csv > gpurun_out/${tag}_ptrace_src.csv 2>/dev/null
python tools/ncu_lines.py gpurun_out/${tag}_ptrace_src.csv 30 > gpurun_out/${tag}_ptrace_lines.txt 2>/dev/null
rm -f gpurun_out/${tag}_ptrace_src.csv
head -14 gpurun_out/${tag}_ptrace_lines.txt
