cd $GRAFT_REPO_ROOT
for cfg in "128 1024" "128 512" "128 256" "256 512" "64 512"; do
  set -- $cfg
  MPL_EXTRA_NVCC_FLAGS="-DMPL_SCORE_THREADS=$1 -DMPL_SCORE_CHUNK=$2" python -c "import __graft_entry__ as g; g.build()" > /dev/null 2>&1
  TAG="threads=$1 chunk=$2" timeout 200 python tools/c3_time.py 2>&1 | tail -1
done
