#!/usr/bin/env python
"""Prints the per-stage table of a bench.py JSON line."""
import json, sys
j = json.load(open(sys.argv[1]))
print(f"value {j['value']:.1f} {j['unit']}  ms/step {j['ms_per_step']:.3f}  fwd-only {j.get('fwd_only_ms_per_frame', 0):.3f} ms  "
      f"e2e {j['e2e']['value']:.1f}  launches {j.get('gpu_launches')}")
for k, v in j.get("stages", {}).items():
    extra = f" fp32 {v['fp32_frac']:.2f} roofline {v['roofline_frac']:.2f}" if "fp32_frac" in v else ""
    print(f"  {k:16s} {v['ms'] * 1e3:8.1f} us  {v.get('GBps', 0):7.0f} GB/s  hbm {v.get('hbm_frac', 0):.2f}{extra}")
print("  scene", j.get("scene"))
