set -x
mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_gpu_host_api.py -x -q -m gpu -rs > gpurun_out/r03_gpu_tests_host_api.log 2>&1; echo "rc=$?" >> gpurun_out/r03_gpu_tests_host_api.log )
tail -6 gpurun_out/r03_gpu_tests_host_api.log
