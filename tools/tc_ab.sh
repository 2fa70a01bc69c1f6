python matchinglib_poselib_b200/build.py --force 2>&1 | grep -i "error"
timeout 600 python -m pytest tests/test_knn_gpu.py tests/test_sharded_gpu.py -m gpu -q -x -k "l2 or c2 or tc" > gpurun_out/pytest_tc.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_tc.log
TAG=full python tools/tc_time.py 2>&1 | tail -1
TAG=full_signed python tools/tc_time.py signed 2>&1 | tail -1
for mode in 1 2; do
  MPL_EXTRA_NVCC_FLAGS="-DMPL_TC_DEBUG=$mode" python matchinglib_poselib_b200/build.py --force 2>&1 | grep -i "error"
  TAG="debug$mode" python tools/tc_time.py 2>&1 | tail -1
done
python matchinglib_poselib_b200/build.py --force 2>&1 | grep -i "error"
