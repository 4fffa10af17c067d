set -x
mkdir -p gpurun_out
( timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/r03_noalloc_alone.log 2>&1; echo "rc=$?" >> gpurun_out/r03_noalloc_alone.log ) 
tail -5 gpurun_out/r03_noalloc_alone.log
( timeout 2400 python -m pytest tests -x -q -m gpu > gpurun_out/r03_gpu_tests_all.log 2>&1; echo "rc=$?" >> gpurun_out/r03_gpu_tests_all.log )
tail -5 gpurun_out/r03_gpu_tests_all.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/r03_bench_n1_v1.json 2> gpurun_out/r03_bench_n1_v1.err; tail -c 1500 gpurun_out/r03_bench_n1_v1.json
