#!/bin/bash
# which bench setting makes the in-bench e2e call slow?  (diagnostic, GPU box)
run() { python bench.py --steps 20 --warmup 3 --no-cpu 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$1', 'e2e ms %.3f' % d['e2e']['ms'], 'step %.4f' % d['ms_per_step'])"; }
run default
MPL_BENCH_NO_GRAPH=1 run nograph
CUDA_DEVICE_MAX_CONNECTIONS=8 run conn8
CUDA_DEVICE_MAX_CONNECTIONS=8 MPL_BENCH_NO_GRAPH=1 run conn8_nograph
run default_again
