#!/bin/bash
# Round-end check on a GPU box: the GPU test tier, smoke(), both bench arms. Outputs under gpurun_out/.
mkdir -p gpurun_out
(time timeout 480 python -m pytest tests -m gpu -x -q) > gpurun_out/final_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/final_tests.log
timeout 60 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/final_smoke.log 2>&1
timeout 400 python bench.py > gpurun_out/final_bench.json 2> gpurun_out/final_bench.err
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/final_ref.json 2> gpurun_out/final_ref.err
tail -3 gpurun_out/final_tests.log
tail -2 gpurun_out/final_smoke.log
cut -c1-300 gpurun_out/final_bench.json
