#!/bin/bash
# One gpurun call: GPU parity tests, the bench line, then ncu (launch list + full capture of the two hot kernels).
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
nproc > gpurun_out/nproc.txt
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?" >> gpurun_out/bench.err
timeout 600 python bench.py --impl reference --steps 30 --warmup 3 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
if [ "$1" = "ncu" ]; then
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/plain.log 2>&1 &&
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/launches.csv python bench.py --steps 5 --warmup 3 > gpurun_out/ncu_launches.log 2>&1 &&
timeout 600 python tools/profile_hot.py > gpurun_out/plain_hot.log 2>&1 &&
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:'scan_kernel|align_kernel|nn_residual|shard_' -c 45 -o gpurun_out/prof python tools/profile_hot.py > gpurun_out/ncu_full.log 2>&1
echo "ncu rc=$?" >> gpurun_out/ncu_full.log
# the summary is made here as well: a report above ~45 MB would push gpurun_out/ over its 64 MiB limit and is dropped
python tools/ncu_summary.py gpurun_out/prof.ncu-rep gpurun_out/ncu_summary >> gpurun_out/ncu_full.log 2>&1
python tools/ncu_launch_summary.py gpurun_out/launches.csv gpurun_out/ncu_launches_summary.txt >> gpurun_out/ncu_full.log 2>&1
if [ "$(stat -c %s gpurun_out/prof.ncu-rep 2>/dev/null || echo 0)" -gt 45000000 ]; then rm -f gpurun_out/prof.ncu-rep; echo "report dropped (size)" >> gpurun_out/ncu_full.log; fi
fi
tail -5 gpurun_out/pytest_gpu.log; cat gpurun_out/bench.json; tail -3 gpurun_out/bench.err
