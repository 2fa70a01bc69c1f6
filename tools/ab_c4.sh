#!/bin/bash
# A/B of prebuilt library variants (ab_libs/*.so, built here with MPL_EXTRA_NVCC_FLAGS) on the C4 pair-batch pipeline
cp matchinglib_poselib_b200/lib/libmplb200.so /tmp/lib_keep.so
for f in ab_libs/lib_*.so; do
  cp $f matchinglib_poselib_b200/lib/libmplb200.so; touch matchinglib_poselib_b200/lib/libmplb200.so
  echo $(basename $f); timeout 200 python tools/c4_sizes.py 4096 4096 2>&1 | tail -2 | cut -c1-150
  timeout 100 python tools/c4_prof.py 512 2>&1 | tail -1 | cut -c1-260
done
cp /tmp/lib_keep.so matchinglib_poselib_b200/lib/libmplb200.so
