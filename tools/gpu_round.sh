#!/bin/bash
# One GPU visit: parity tests, smoke, bench (+ optional ncu passes: NCU=1; reference arm: REF=1).
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -s ${PYTEST_ARGS} > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -8 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/smoke.log
if [ -z "$NO_BENCH" ]; then
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
head -c 1500 gpurun_out/bench.json; tail -5 gpurun_out/bench.err
fi
if [ -n "$NCU" ]; then
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu_launch.log 2>&1; echo "ncu launches rc=$?"
# kernel names are matched on the demangled name (template arguments included); NCU_KERNELS overrides the list
# the ranking kernel in Hamming mode: the tool runs popc, then the bf16 form (13 chunks x 3 repetitions = launches 0..38), then e4m3
# (7 chunks x 3): launch 41 is a full 196 608-row e4m3 chunk of the 10 000 x 1.25 M scan
timeout 600 ncu --set full --clock-control none --import-source on -k "regex:l2_tc_rank_kernel" -s 41 -c 1 -f -o gpurun_out/full_l2_tc_rank_kernel_hamming python tools/hamming_tc_time.py 10000 1250000 32 > gpurun_out/ncu_full_tc_hamming.log 2>&1; echo "ncu full tc hamming rc=$?"
for k in ${NCU_KERNELS:-lmeds_median_kernel hamming_prep_kernel}; do
  f=$(echo $k | tr -d '<>')
  timeout 600 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k "regex:$k" -s 1 -c 1 -f -o gpurun_out/full_$f python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu_full_$f.log 2>&1; echo "ncu full $k rc=$?"
done
# the tf32 mode of the ranking kernel: the 3rd .. launches of l2_tc_rank_kernel belong to parts.l2_tf32_c2 (after the C2 headline's)
[ -n "$NCU_TF32" ] && timeout 600 ncu --set full --clock-control none --import-source on -k "regex:l2_tc_rank_kernel" -s 40 -c 1 -f -o gpurun_out/full_l2_tc_rank_kernel_tf32 python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu_full_tc_tf32.log 2>&1; echo "ncu full tf32 rc=$?"
fi
if [ -n "$REF" ]; then
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "reference arm rc=$?"
cat gpurun_out/bench_reference.json
fi
