#!/bin/bash
# round 2, 2-GPU call: the e2e experiment tool at a small size (smoke test before the 8-GPU call), ncu counters of the fused
# compute + NVLink kernel
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 tools/e2e_multi.py --size 32768 --variants default,ring,p1=1 --passes 2 --out gpurun_out/c3_e2e_multi_n2.jsonl --timeline-dir gpurun_out > gpurun_out/c3_e2e.log 2>&1; echo "e2e tool rc=$?"
tail -n 12 gpurun_out/c3_e2e.log | cut -c 1-1500
timeout 300 python tools/ncu_fused_fetch.py > gpurun_out/c3_fused_fetch_events.json 2> gpurun_out/c3_fused_fetch.err && \
timeout 600 ncu --metrics gpu__time_duration.sum,nvlrx__bytes.sum,nvltx__bytes.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum,smsp__pipe_tensor_subpipe_dmma_cycles_active.avg,sm__cycles_elapsed.avg,smsp__cycles_active.avg,lts__t_sector_hit_rate.pct --clock-control none -k regex:dgemm_tma_ws_kernel --csv --log-file gpurun_out/c3_ncu_fused_fetch.csv python tools/ncu_fused_fetch.py > gpurun_out/c3_ncu_run.log 2>&1
echo "ncu rc=$?"
cat gpurun_out/c3_fused_fetch_events.json; tail -n 3 gpurun_out/c3_fused_fetch.err
tail -n 5 gpurun_out/c3_ncu_run.log
