#!/bin/bash
# round 2, multi-GPU call: tests/test_gpu_multi.py and bench.py at N = number of GPUs given ($1, default 2)
cd "$(dirname "$0")/.." || exit 1
N=${1:-2}
TAG=${2:-c2}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_multi.py -x -q > gpurun_out/${TAG}_multi_tests.log 2>&1; echo "multi tests rc=$?" | tee -a gpurun_out/${TAG}_summary.txt
BENCH_TIMELINE_FILE=gpurun_out/${TAG}_timeline_n${N}.json timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps ${STEPS:-2} --warmup ${WARMUP:-1} > gpurun_out/${TAG}_bench_n${N}.json 2> gpurun_out/${TAG}_bench_n${N}.err; echo "bench rc=$?" | tee -a gpurun_out/${TAG}_summary.txt
tail -n 5 gpurun_out/${TAG}_multi_tests.log
tail -n 20 gpurun_out/${TAG}_bench_n${N}.err
cat gpurun_out/${TAG}_bench_n${N}.json
