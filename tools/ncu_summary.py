"""Summarises an ncu --set full report (CSV raw page) into a small JSON for profiles/.
    ncu -i rep.ncu-rep --page raw --csv > raw.csv ; python tools/ncu_summary.py raw.csv out.json [size_key traffic.json]"""
import csv, json, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
out = []
KEYS = ['Kernel Name', 'Grid Size', 'Block Size', 'gpu__time_duration.sum', 'launch__registers_per_thread',
        'launch__shared_mem_per_block_dynamic',
        'TPC.TriageCompute.sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed',
        'TPC.TriageCompute.sm__pipe_fp64_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed', 'lts__t_sector_hit_rate.pct',
        'lts__t_bytes.sum', 'smsp__warps_active.avg.per_cycle_active', 'smsp__inst_executed.sum',
        'sm__cycles_elapsed.avg', 'smsp__cycles_active.avg',
        'SM_C.TriageCompute.smsp__pipe_tensor_subpipe_dmma_cycles_active.avg', 'TPC.TriageCompute.sm__cycles_active.avg',
        'dram__bytes.sum.per_second']
for vals in rows[2:]:
    d = {h: (v, u) for h, u, v in zip(hdr, units, vals)}
    o = {k: {"value": d[k][0], "unit": d[k][1]} for k in KEYS if k in d}
    for h in hdr:
        if 'smsp__average_warps_issue_stalled' in h and h.endswith('_per_issue_active.ratio') and float(d[h][0] or 0) > 0.005:
            o[h] = {"value": d[h][0], "unit": d[h][1]}
    out.append(o)
json.dump(out if len(out) > 1 else out[0], open(sys.argv[2], "w"), indent=1)
o = out[0]
def gb(k):
    v, u = float(o[k]["value"]), o[k]["unit"]
    return v * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1, "Tbyte": 1e12}[u]
traffic = gb('dram__bytes_read.sum') + gb('dram__bytes_write.sum')
print("traffic bytes", traffic)
for k in KEYS[3:12]:
    if k in o: print(k, o[k])
if len(sys.argv) > 4:
    try: t = json.load(open(sys.argv[4]))
    except Exception: t = {}
    t[sys.argv[3]] = traffic
    json.dump(t, open(sys.argv[4], "w"), indent=1)
