#!/bin/bash
# One warm nn_residual_tile_kernel launch (262 144 queries, 12 M-point map >> L2) under `ncu --set full` with source
# correlation: details page (dram__throughput, lts__t_sector_hit_rate, sm__warps_active ...) and stall samples by line.
mkdir -p gpurun_out
timeout 300 python tools/bench_nn.py > gpurun_out/nn1_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'^nn_residual_tile_kernel$' -s 6 -c 1 -o gpurun_out/nn1 python tools/bench_nn.py > gpurun_out/nn1_ncu.log 2>&1
ncu -i gpurun_out/nn1.ncu-rep --page source --csv > gpurun_out/nn1_source.csv 2>> gpurun_out/nn1_ncu.log
python tools/ncu_hot_lines.py gpurun_out/nn1_source.csv gpurun_out/nn1_hot_lines.txt 60 >> gpurun_out/nn1_ncu.log 2>&1
ncu -i gpurun_out/nn1.ncu-rep --page details > gpurun_out/nn1_details.txt 2>> gpurun_out/nn1_ncu.log
ncu -i gpurun_out/nn1.ncu-rep --page raw --csv > gpurun_out/nn1_raw.csv 2>> gpurun_out/nn1_ncu.log
head -40 gpurun_out/nn1_hot_lines.txt
rm -f gpurun_out/nn1_source.csv
if [ "$(stat -c %s gpurun_out/nn1.ncu-rep 2>/dev/null || echo 0)" -gt 20000000 ]; then rm -f gpurun_out/nn1.ncu-rep; fi
