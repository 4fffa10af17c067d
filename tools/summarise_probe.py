"""Turns peak_probe's JSON lines into profiles/fp64_peak_probe.json (read by bench.py for roofline.dmma_probe_tflops)."""
import json, sys
rows = [json.loads(l) for l in open(sys.argv[1]) if l.startswith("{")]
dev = rows[0]
dmma = [r for r in rows if r.get("probe") == "dmma_m8n8k4"]
dfma = [r for r in rows if r.get("probe") == "dfma"]
best_dmma = max(dmma, key=lambda r: r["flop_per_clk_per_sm"])
best_dfma = max(dfma, key=lambda r: r["flop_per_clk_per_sm"])
out = {"device": dev, "dmma_tflops": best_dmma["tflops"], "dmma_flop_per_clk_per_sm": best_dmma["flop_per_clk_per_sm"],
       "dmma_sm_mhz": best_dmma["sm_mhz_inkernel"], "dmma_latency_clk": min(r["clk_per_instr_per_smsp"] for r in dmma if r["nacc"] == 1),
       "dfma_tflops": best_dfma["tflops"], "dfma_flop_per_clk_per_sm": best_dfma["flop_per_clk_per_sm"], "rows": rows}
json.dump(out, open(sys.argv[2], "w"), indent=1)
print({k: v for k, v in out.items() if k != "rows"})
