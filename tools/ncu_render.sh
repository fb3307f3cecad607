#!/bin/bash
# usage: bash tools/ncu_render.sh <tag> [env...]: full ncu capture of one k_render launch of the short C5 bench
tag=$1; shift
SMALL="python bench.py --steps 2 --warmup 1 --spp-per-step 4 --no-cpu-baseline --no-lens-grad"
env "$@" $SMALL > gpurun_out/${tag}_plain.log 2>&1 &&
env "$@" timeout 1200 ncu --set full --clock-control none --import-source on -k regex:k_render -s 1 -c 1 -f -o gpurun_out/${tag}_render $SMALL > gpurun_out/${tag}_ncu.log 2>&1
echo "rc=$?"; tail -1 gpurun_out/${tag}_plain.log | cut -c1-200
