#!/bin/bash
# A/B of prebuilt library variants (ab_libs/lib_<threads>_<chunk>.so, built here with MPL_EXTRA_NVCC_FLAGS) on the C3 scoring kernel
cp matchinglib_poselib_b200/lib/libmplb200.so /tmp/lib_keep.so
for f in ab_libs/lib_*.so; do
  cp $f matchinglib_poselib_b200/lib/libmplb200.so; touch matchinglib_poselib_b200/lib/libmplb200.so
  TAG=$(basename $f) timeout 200 python tools/c3_time.py 2>&1 | tail -1
done
cp /tmp/lib_keep.so matchinglib_poselib_b200/lib/libmplb200.so
