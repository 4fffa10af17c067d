set -x
mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_gpu_out_of_core.py tests/test_gpu_multi.py tests/test_gpu_rank_pipeline.py tests/test_zz_no_alloc.py -x -q -m gpu -rs > gpurun_out/r03_gpu_tests_subset_v2.log 2>&1; echo "rc=$?" >> gpurun_out/r03_gpu_tests_subset_v2.log )
tail -15 gpurun_out/r03_gpu_tests_subset_v2.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/r03_bench_n1_v2.json 2> gpurun_out/r03_bench_n1_v2.err; echo "bench rc=$?"; tail -3 gpurun_out/r03_bench_n1_v2.err
python - <<'PY'
import json
l=json.loads(open('gpurun_out/r03_bench_n1_v2.json').read().strip().splitlines()[-1])
print(l['value'], l['e2e']['value'] if l.get('e2e') else None)
c=l.get('configs',{})
print(c.get('error'), c.get('seconds'), c.get('max_err_over_bound'))
for r in c.get('shapes',[]):
    print(r['M'],r['N'],r['K'],r['lda'],r['alpha'],r['beta'],r['tile'],'%.2f'%r['kernel_tflops'],'%.3f'%r['frac'],'cublas %.2f'%r.get('cublas_tflops',0),'e2e %.2f'%r['e2e_tflops'],'%.2e'%r['max_err_over_bound'],'%.2e'%r['e2e_max_err_over_bound'], r.get('hbm_gbs'))
PY
