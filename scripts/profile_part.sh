#!/bin/bash
# --set full capture of the binning kernels of one frame (default bench command), after a plain run exited 0.
export FDLS_BENCH_FRAMES_PER_STEP=2 FDLS_BENCH_TRAIN=0
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --workload ${FDLS_PROFILE_WORKLOAD:-cfg2}"
mkdir -p gpurun_out
$CMD > gpurun_out/plain_part.log 2>&1 || { tail -5 gpurun_out/plain_part.log; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:"part_|scan_" -s 10 -c 5 -f -o gpurun_out/r02_part $CMD > gpurun_out/ncu_p.log 2>&1
tail -3 gpurun_out/ncu_p.log; ls -la gpurun_out/*.ncu-rep
