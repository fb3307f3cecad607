#!/bin/bash
# One GPU-box pass: parity tests, bench line, ncu launch list and one full capture of the render kernel.
# usage (under gpurun): bash tools/gpu_check.sh <tag>
tag=${1:-r01}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_gpu_tests.log 2>&1; echo "tests rc=$?"
tail -3 gpurun_out/${tag}_gpu_tests.log
timeout 600 python bench.py > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench rc=$?"
cat gpurun_out/${tag}_bench.json
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${tag}_bench_ref.json 2>> gpurun_out/${tag}_bench.err; echo "ref rc=$?"
cat gpurun_out/${tag}_bench_ref.json
SMALL="python bench.py --steps 2 --warmup 1 --spp-per-step 4 --no-cpu-baseline"
timeout 300 $SMALL > gpurun_out/${tag}_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches.csv $SMALL > gpurun_out/${tag}_ncu1.log 2>&1
echo "launch list rc=$?"
timeout 300 $SMALL > gpurun_out/${tag}_plain2.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:k_render -s 1 -c 1 -f -o gpurun_out/${tag}_render $SMALL > gpurun_out/${tag}_ncu2.log 2>&1
echo "full capture rc=$?"
ls -la gpurun_out
