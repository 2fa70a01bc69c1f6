#!/bin/bash
mkdir -p gpurun_out
bash tools/run_tc.sh
timeout 600 ncu --set full --clock-control none --import-source on -k regex:l2_tc_rank_kernel -s 2 -c 1 -f -o gpurun_out/full_tc3 python bench.py --steps 2 --warmup 1 --no-cpu --no-parts > gpurun_out/ncu_full_tc3.log 2>&1; echo "ncu rc=$?"
