#!/bin/bash
mkdir -p gpurun_out
./tools/microbench > gpurun_out/microbench.json 2> gpurun_out/microbench.err; cat gpurun_out/microbench.json
bash tools/run_tc.sh
timeout 600 ncu --set full --clock-control none --import-source on -k regex:l2_tc_rank_kernel -s 2 -c 1 -f -o gpurun_out/full_tc2 python bench.py --steps 2 --warmup 1 --no-cpu --no-parts > gpurun_out/ncu_full_tc2.log 2>&1; echo "ncu rc=$?"
