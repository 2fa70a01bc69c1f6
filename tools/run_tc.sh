timeout 300 python -m pytest tests/test_knn_gpu.py -x -q -k "tc" 2>&1 | tail -3; python bench.py --steps 10 --warmup 3 --no-cpu --no-parts | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['roofline']['kernel_ms'], d['roofline']['frac'], d['roofline']['step_share'], d['e2e']['ms'])"
