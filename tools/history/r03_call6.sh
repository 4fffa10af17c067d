set -x
mkdir -p gpurun_out
free -g | head -2
timeout 2400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29518 bench.py --gpus 8 --steps 5 --warmup 3 > gpurun_out/r03_bench_n8_v1.json 2> gpurun_out/r03_bench_n8_v1.err; echo "bench rc=$?"; tail -5 gpurun_out/r03_bench_n8_v1.err
python - <<'PY'
import json
l=json.loads(open('gpurun_out/r03_bench_n8_v1.json').read().strip().splitlines()[-1])
print(l['value'], l['parity_blocks_max_ratio'], l['e2e']['value'], l['e2e']['parity_blocks_max_ratio'], l['e2e']['parity_blocks'])
print(json.dumps(l['parity_panels']['cases']))
v=l.get('e2e_v21'); print(v and (v['value'], v['size'], v['ms_per_step'], v['spot_check_max_err_over_bound'], [(o['value'],o['size']) for o in v.get('other_sizes',[])]))
PY
