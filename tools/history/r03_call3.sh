set -x
mkdir -p gpurun_out
( timeout 600 python -m pytest tests/test_gpu_host_api.py -x -q -m gpu -rs > gpurun_out/r03_gpu_tests_host_api.log 2>&1; echo "rc=$?" >> gpurun_out/r03_gpu_tests_host_api.log )
tail -6 gpurun_out/r03_gpu_tests_host_api.log
timeout 900 python bench.py --steps 5 --warmup 3 --no-configs > gpurun_out/r03_bench_n1_v3.json 2> gpurun_out/r03_bench_n1_v3.err; echo "bench rc=$?"; tail -3 gpurun_out/r03_bench_n1_v3.err
python - <<'PY'
import json
l=json.loads(open('gpurun_out/r03_bench_n1_v3.json').read().strip().splitlines()[-1])
print(l['value'], l['parity_blocks_max_ratio'], l['e2e']['value'], l['e2e']['parity_blocks_max_ratio'], l['e2e']['parity_blocks'])
PY
