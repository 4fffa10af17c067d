set -x
mkdir -p gpurun_out
( timeout 2400 python -m pytest tests -x -q -m gpu -rs > gpurun_out/r03_gpu_tests_1gpu_v3.log 2>&1; echo "rc=$?" >> gpurun_out/r03_gpu_tests_1gpu_v3.log )
tail -12 gpurun_out/r03_gpu_tests_1gpu_v3.log
timeout 600 python bench.py --steps 3 --warmup 3 --no-scale-base --no-cpu-baseline --no-configs > gpurun_out/r03_bench_n1_v5.json 2>/dev/null; python -c "
import json; l=json.loads(open('gpurun_out/r03_bench_n1_v5.json').read().strip().splitlines()[-1]); print(l['value'], l['roofline']['traffic'], l['roofline']['traffic_source'][:80])"
