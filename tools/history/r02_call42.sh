#!/bin/bash
# round 2, 1-GPU call, final build: ncu launch list of the bench command; ncu --set full of the kernel at 16384^3 and at 1024^3
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
O=gpurun_out
timeout 300 python bench.py --steps 2 --warmup 3 --no-scale-base --no-cpu-baseline > $O/c42_bench_plain.json 2> $O/c42_bench_plain.err; echo "bench rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/c42_ncu_launches_bench.csv python bench.py --steps 2 --warmup 3 --no-scale-base --no-cpu-baseline > $O/c42_ncu_bench.log 2>&1; echo "ncu launches rc=$?"
for shape in "16384 16384 16384" "1024 1024 1024"; do
  tag=$(echo $shape | tr ' ' 'x')
  timeout 120 python tools/one_shape.py $shape > $O/c42_plain_$tag.json || exit 1
  timeout 400 ncu --set full --clock-control none --import-source on -k regex:dgemm_tma -s 2 -c 1 -o $O/c42_full_$tag -f python tools/one_shape.py $shape > $O/c42_ncu_$tag.log 2>&1
  ncu -i $O/c42_full_$tag.ncu-rep --page raw --csv > $O/c42_full_$tag.raw.csv 2>/dev/null
done
python - <<'P'
import csv, collections
rows = [r for r in csv.reader(open("gpurun_out/c42_ncu_launches_bench.csv")) if len(r) > 5]
hdr = rows[0]; k = hdr.index("Kernel Name"); v = hdr.index("Metric Value")
t = collections.Counter(); n = collections.Counter()
for r in rows[1:]:
    name = r[k].split("(")[0]; t[name] += float(r[v].replace(",", "")); n[name] += 1
tot = sum(t.values())
for name, x in t.most_common(8): print(f"{100*x/tot:6.2f} %  {n[name]:4d} launches  {name}")
P
ls -la $O | grep c42
