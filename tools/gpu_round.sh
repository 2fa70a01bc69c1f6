#!/bin/bash
# One GPU visit: parity tests, smoke, bench (+ optional ncu passes: NCU=1; reference arm: REF=1).
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -s ${PYTEST_ARGS} > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -25 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/smoke.log
if [ -z "$NO_BENCH" ]; then
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
cat gpurun_out/bench.json; tail -5 gpurun_out/bench.err
fi
if [ -n "$NCU" ]; then
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 1 --no-cpu > gpurun_out/ncu_launch.log 2>&1; echo "ncu launches rc=$?"
for k in l2_tc_rank_kernel hamming_top2_kernel sampson_score_kernel fivept_solve_kernel; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$k -s 2 -c 1 -f -o gpurun_out/full_$k python bench.py --steps 2 --warmup 1 --no-cpu > gpurun_out/ncu_full_$k.log 2>&1; echo "ncu full $k rc=$?"
done
fi
if [ -n "$REF" ]; then
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "reference arm rc=$?"
cat gpurun_out/bench_reference.json
fi
