#!/bin/bash
# round 2, first GPU call: new rank-pipeline tests (emulated peers), whole -m gpu suite, A/B of the round-1 library against
# this build, a short bench run
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/c1_smi.txt 2>&1
timeout 600 python -m pytest tests/test_gpu_rank_pipeline.py -x -q > gpurun_out/c1_rank_tests.log 2>&1; echo "rank tests rc=$?" | tee -a gpurun_out/c1_summary.txt
P=optimizing-dgemm-on-intel-cpus-with-avx512f_b200
for shape in "16384 16384 16384 3" "4096 4096 4096 10" "2048 2048 2048 20" "1024 1024 1024 30" "512 512 512 50" "1000 1003 997 30" "512 512 16384 20"; do
  for lib in libdgemm_b200_r1.so libdgemm_b200.so libdgemm_b200_r1.so libdgemm_b200.so; do
    timeout 300 python tools/ab_lib.py $P/$lib $shape >> gpurun_out/c1_ab.jsonl 2>> gpurun_out/c1_ab.err
  done
done
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/c1_gpu_tests.log 2>&1; echo "gpu tests rc=$?" | tee -a gpurun_out/c1_summary.txt
timeout 600 python bench.py --steps 3 --warmup 3 > gpurun_out/c1_bench.json 2> gpurun_out/c1_bench.err; echo "bench rc=$?" | tee -a gpurun_out/c1_summary.txt
tail -3 gpurun_out/c1_rank_tests.log gpurun_out/c1_gpu_tests.log
cat gpurun_out/c1_ab.jsonl
