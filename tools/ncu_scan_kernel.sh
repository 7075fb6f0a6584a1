#!/bin/bash
# One warm scan_kernel launch under `ncu --set full` with source correlation; summaries into gpurun_out/ (the report itself only if small).
mkdir -p gpurun_out
export PYTHONPATH="$(pwd)${PYTHONPATH:+:$PYTHONPATH}"
cat > /tmp/drive_scans.py <<'PY'
import lidar_odometry_demo_b200 as lo
from bench import make_stream
ctx = lo.Context(0); od = lo.LidarOdometry(ctx)
for sc in make_stream(24, 0):
    od.processCloud(sc)
print(od.getCurrentPose())
PY
timeout 300 python /tmp/drive_scans.py > gpurun_out/scan1_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'^scan_kernel$' -s 20 -c 1 -o gpurun_out/scan1 python /tmp/drive_scans.py > gpurun_out/scan1_ncu.log 2>&1
ncu -i gpurun_out/scan1.ncu-rep --page source --csv > gpurun_out/scan1_source.csv 2>> gpurun_out/scan1_ncu.log
python tools/ncu_hot_lines.py gpurun_out/scan1_source.csv gpurun_out/scan1_hot_lines.txt 60 >> gpurun_out/scan1_ncu.log 2>&1
ncu -i gpurun_out/scan1.ncu-rep --page details > gpurun_out/scan1_details.txt 2>> gpurun_out/scan1_ncu.log
ls -la gpurun_out/scan1* ; head -70 gpurun_out/scan1_hot_lines.txt
if [ "$(stat -c %s gpurun_out/scan1_source.csv 2>/dev/null || echo 0)" -gt 20000000 ]; then rm -f gpurun_out/scan1_source.csv; fi
if [ "$(stat -c %s gpurun_out/scan1.ncu-rep 2>/dev/null || echo 0)" -gt 30000000 ]; then rm -f gpurun_out/scan1.ncu-rep; fi
