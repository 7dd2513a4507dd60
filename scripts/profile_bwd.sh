#!/bin/bash
# One --set full capture of render_bwd / preprocess_bwd of the cfg3 train step (default bench command), after a plain run.
export FDLS_BENCH_FRAMES_PER_STEP=2
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
mkdir -p gpurun_out
$CMD > gpurun_out/plain_bwd.log 2>&1 || { tail -5 gpurun_out/plain_bwd.log; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:"render_bwd|preprocess_bwd" -s 4 -c 2 -f -o gpurun_out/r02_bwd $CMD > gpurun_out/ncu_b.log 2>&1
tail -3 gpurun_out/ncu_b.log; ls -la gpurun_out/*.ncu-rep
