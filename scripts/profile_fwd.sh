#!/bin/bash
# --set full capture of the forward compositor and preprocess of one frame (default bench command), after a plain run.
export FDLS_BENCH_FRAMES_PER_STEP=2 FDLS_BENCH_TRAIN=0
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
mkdir -p gpurun_out
$CMD > gpurun_out/plain_fwd.log 2>&1 || { tail -5 gpurun_out/plain_fwd.log; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:"render_fwd|preprocess_fwd" -s 4 -c 2 -f -o gpurun_out/r02_fwd $CMD > gpurun_out/ncu_f.log 2>&1
tail -3 gpurun_out/ncu_f.log; ls -la gpurun_out/*.ncu-rep
