#!/bin/bash
# Does NVML polling inside bench.py cause the multi-millisecond host stalls seen on some boxes? Alternates runs with and
# without the clock sampler on ONE box and prints the per-pass step times (findings: DESIGN.md section 7).
mkdir -p gpurun_out
for r in 1 2 3; do
  for mode in clk noclk; do
    if [ $mode = noclk ]; then export LO_BENCH_NO_CLOCKS=1; else unset LO_BENCH_NO_CLOCKS; fi
    timeout 300 python bench.py > gpurun_out/exp_${mode}_$r.json 2>/dev/null
    python - <<EOF
import json
d=json.load(open("gpurun_out/exp_${mode}_$r.json"))
a=d["attempts"]; print("$mode", $r, round(d["value"]), [round(x,3) for x in a["resident_ms_per_step"]], [round(x,3) for x in a["e2e_ms_per_step"]], a["step_ms_percentiles"]["max"], d["one_call"]["value"], d["clocks"]["samples"])
EOF
  done
done
