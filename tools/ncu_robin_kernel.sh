#!/bin/bash
# One robin_order_block_kernel launch (n = 13 000 random hashes) under `ncu --set full` with source correlation.
mkdir -p gpurun_out
export PYTHONPATH="$(pwd)${PYTHONPATH:+:$PYTHONPATH}"
cat > /tmp/drive_robin.py <<'PY'
import numpy as np, lidar_odometry_demo_b200 as lo
ctx = lo.Context(0)
h = np.random.default_rng(5).choice(1 << 22, 13000, replace=False).astype(np.uint32)
for _ in range(3):
    ctx.robin_iteration_order(h)
PY
timeout 300 ncu --set full --clock-control none --import-source on -k regex:'robin_order_block_kernel' -s 2 -c 1 -o gpurun_out/robin1 python /tmp/drive_robin.py > gpurun_out/robin1_ncu.log 2>&1
ncu -i gpurun_out/robin1.ncu-rep --page source --csv > gpurun_out/robin1_source.csv 2>> gpurun_out/robin1_ncu.log
python tools/ncu_hot_lines.py gpurun_out/robin1_source.csv gpurun_out/robin1_hot_lines.txt 60 >> gpurun_out/robin1_ncu.log 2>&1
ncu -i gpurun_out/robin1.ncu-rep --page details > gpurun_out/robin1_details.txt 2>> gpurun_out/robin1_ncu.log
head -75 gpurun_out/robin1_hot_lines.txt
rm -f gpurun_out/robin1_source.csv
if [ "$(stat -c %s gpurun_out/robin1.ncu-rep 2>/dev/null || echo 0)" -gt 20000000 ]; then rm -f gpurun_out/robin1.ncu-rep; fi
