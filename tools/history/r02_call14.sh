#!/bin/bash
# round 2, 1-GPU call: whole -m gpu suite on the current tree, default bench, reference arm, sweep, ncu launch list of the bench command
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
O=gpurun_out
( time timeout 1500 python -m pytest tests -m gpu -x -q ) > $O/c14_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 6 $O/c14_pytest.log
timeout 900 python bench.py > $O/c14_bench_n1.json 2> $O/c14_bench_n1.err; echo "bench rc=$?"; tail -n 3 $O/c14_bench_n1.err; cut -c1-1500 $O/c14_bench_n1.json
timeout 900 python tools/sweep.py > $O/c14_sweep.jsonl 2> $O/c14_sweep.err; echo "sweep rc=$?"
python - <<'PY'
import json
for l in open("gpurun_out/c14_sweep.jsonl"):
    r = json.loads(l); print(r["M"], r["N"], r["K"], r.get("label", "")[:40], "ours", r["kernel_tflops"], "cublas", r.get("cublas_tflops"), "e2e", r.get("e2e_tflops"))
PY
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/c14_ncu_launches_bench.csv python bench.py --steps 2 --warmup 3 --no-scale-base --no-cpu-baseline > $O/c14_ncu_bench.log 2>&1; echo "ncu launches rc=$?"
python - <<'PY'
import csv, collections
rows = [r for r in csv.reader(open("gpurun_out/c14_ncu_launches_bench.csv")) if len(r) > 5]
hdr = rows[0]; ik = hdr.index("Kernel Name"); iv = hdr.index("Metric Value"); iu = hdr.index("Metric Unit")
tot = collections.defaultdict(float); cnt = collections.Counter()
for r in rows[1:]:
    try: v = float(r[iv].replace(",", ""))
    except ValueError: continue
    v *= {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(r[iu], 1e-6)
    tot[r[ik][:70]] += v; cnt[r[ik][:70]] += 1
s = sum(tot.values())
for k, v in sorted(tot.items(), key=lambda kv: -kv[1])[:8]: print(f"{v:10.2f} ms {100*v/s:6.2f} % n={cnt[k]:4d} {k}")
PY
