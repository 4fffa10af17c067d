set -x
mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_gpu_host_api.py tests/test_gpu_multi.py -x -q -m gpu -rs > gpurun_out/r03_gpu_tests_2gpu.log 2>&1; echo "rc=$?" >> gpurun_out/r03_gpu_tests_2gpu.log )
tail -8 gpurun_out/r03_gpu_tests_2gpu.log
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 3 --warmup 3 --size 16384 > gpurun_out/r03_bench_n2_16384_v1.json 2> gpurun_out/r03_bench_n2_16384_v1.err; echo "bench rc=$?"; tail -5 gpurun_out/r03_bench_n2_16384_v1.err
python - <<'PY'
import json
l=json.loads(open('gpurun_out/r03_bench_n2_16384_v1.json').read().strip().splitlines()[-1])
print(l['value'], l['parity_blocks_max_ratio'], l['e2e']['value'], l['e2e']['parity_blocks_max_ratio'], l['e2e']['parity_blocks'])
print(json.dumps(l['parity_panels']['cases']))
v=l.get('e2e_v21'); print(v and (v['value'], v['size'], v['spot_check_max_err_over_bound']))
PY
