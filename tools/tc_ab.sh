for lag in 0 600 1200; do
  if [ $lag = 0 ]; then FL=""; else FL="-DMPL_TC_LAG=$lag"; fi
  MPL_EXTRA_NVCC_FLAGS="$FL" timeout 300 python matchinglib_poselib_b200/build.py --force 2>&1 | grep -i "error"
  TAG="lag$lag" timeout 100 python tools/tc_time.py 2>&1 | tail -1
  MPL_EXTRA_NVCC_FLAGS="$FL -DMPL_TC_DEBUG=1" timeout 300 python matchinglib_poselib_b200/build.py --force 2>&1 | grep -i "error"
  TAG="lag$lag-mmaonly" timeout 100 python tools/tc_time.py 2>&1 | tail -1
done
timeout 300 python matchinglib_poselib_b200/build.py --force 2>&1 | grep -i "error"
