#!/usr/bin/env python
"""Human-readable digest of a bench.py / kernels_json.py JSON (last line of the file)."""
import json
import sys

txt = open(sys.argv[1]).read().strip()
try:
    j = json.loads(txt)
except Exception:
    j = json.loads(txt.splitlines()[-1])
if "ms_per_step" in j:
    e = j.get("e2e", {})
    print(f"n_gpus {j.get('n_gpus')}  ms/step {j['ms_per_step']:.2f}  steps/s {j['value']:.3f}  step TF/s {j.get('step_tflops', 0):.0f} "
          f"(nominal frac {j.get('step_frac_of_nominal_peak', 0):.3f})  clocks {j.get('clocks')}")
    print(f"e2e: {e.get('value', 0):.3f} steps/s, call {e.get('call_s', 0):.3f} s, first call {e.get('first_call_s', 0):.3f} s, setup {e.get('setup_ms', 0):.0f} ms; "
          f"s/image {j.get('s_per_image', 0):.3f}; launches/step {j.get('gpu_launches_per_step')}")
    print("parity:", j.get("parity"))
    if j.get("invalid"):
        print("INVALID:", j["invalid"])
    if "cpu_baseline" in j:
        print("cpu_baseline:", {k: v for k, v in j["cpu_baseline"].items() if k != "sample"})
rw = j.get("roofline_worst")
if rw:
    print(f"roofline_worst: {rw['kernel']} {rw['achieved']:.0f} TF/s of {rw['peak']:.0f} = {rw['frac']:.3f}; launch {rw['launch_us']:.1f} us; share of block {rw['share_of_block']:.3f}")
r = j.get("roofline")
if r:
    print(f"roofline: {r['achieved']:.0f} TF/s of {r['peak']:.0f} = {r['frac']:.3f}; launch {r['launch_us']:.1f} us; cublas {r.get('cublas_same_shape_tflops')}; traffic {r.get('traffic')} vs alg {r.get('algorithmic_bytes')}")
for k, v in j.get("kernels", {}).items():
    print(" ", k, json.dumps(v)[:330] if isinstance(v, dict) else v)
