#!/bin/bash
# round 3, 1-GPU call: full GPU suite on the final build; bench N=1; ncu launch list of the bench command; ncu --set full at 16384^3
set -x
mkdir -p gpurun_out
O=gpurun_out
( timeout 1500 python -m pytest tests -x -q -m gpu -rs > $O/r03_gpu_tests_1gpu_v2.log 2>&1; echo "rc=$?" >> $O/r03_gpu_tests_1gpu_v2.log )
tail -12 $O/r03_gpu_tests_1gpu_v2.log
timeout 900 python bench.py > $O/r03_bench_n1_v4.json 2> $O/r03_bench_n1_v4.err; echo "bench rc=$?"
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > $O/r03_bench_reference_arm.json 2> $O/r03_bench_reference_arm.err; echo "ref arm rc=$?"; tail -c 600 $O/r03_bench_reference_arm.json
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file $O/r03_ncu_launches_bench.csv python bench.py --steps 2 --warmup 3 --no-scale-base --no-cpu-baseline > $O/r03_ncu_bench.log 2>&1; echo "ncu launches rc=$?"
timeout 120 python tools/one_shape.py 16384 16384 16384 > $O/r03_plain_16384.json || exit 1
timeout 500 ncu --set full --clock-control none --import-source on -k regex:dgemm_tma -s 2 -c 1 -o $O/r03_full_16384 -f python tools/one_shape.py 16384 16384 16384 > $O/r03_ncu_16384.log 2>&1
ncu -i $O/r03_full_16384.ncu-rep --page raw --csv > $O/r03_full_16384.raw.csv 2>/dev/null
rm -f $O/r03_full_16384.ncu-rep
python - <<'P'
import csv, collections, json
rows = [r for r in csv.reader(open("gpurun_out/r03_ncu_launches_bench.csv")) if len(r) > 5]
hdr = rows[0]; k = hdr.index("Kernel Name"); v = hdr.index("Metric Value")
t = collections.Counter(); n = collections.Counter()
for r in rows[1:]:
    name = r[k].split("(")[0]; t[name] += float(r[v].replace(",", "")); n[name] += 1
tot = sum(t.values())
for name, x in t.most_common(8): print(f"{100*x/tot:6.2f} %  {n[name]:4d} launches  {name}")
l=json.loads(open('gpurun_out/r03_bench_n1_v4.json').read().strip().splitlines()[-1])
print(l['value'], l['e2e']['value'], l['roofline']['frac'], l['roofline']['traffic_source'][:60], l['configs']['max_err_over_bound'])
P
