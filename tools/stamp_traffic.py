"""Stamps profiles/ncu_traffic.json with the fingerprint of the kernel sources its figures were captured on, so that bench.py
can tell a current roofline.traffic from a stale one (bench.kernel_sources_sha16).  Run right after tools/ncu_summary.py has
written the new figures from an `ncu --set full` capture of THIS build:
    python tools/stamp_traffic.py"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
p = os.path.join(ROOT, "profiles", "ncu_traffic.json")
t = json.load(open(p))
t["kernel_sources_sha16"] = bench.kernel_sources_sha16()
json.dump(t, open(p, "w"), indent=1)
print(t["kernel_sources_sha16"])
