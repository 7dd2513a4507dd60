#!/bin/bash
# One --set full capture of the fused deformation MLP kernel (bench.py --mode deform), after a plain run exited 0.
export FDLS_BENCH_FRAMES_PER_STEP=2
CMD="python bench.py --mode deform --steps 1 --warmup 3 --no-cpu-baseline"
mkdir -p gpurun_out
$CMD > gpurun_out/plain_deform.log 2>&1 || { tail -5 gpurun_out/plain_deform.log; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:"deform_mlp_kernel" -s 3 -c 1 -f -o gpurun_out/r02_deform_mlp $CMD > gpurun_out/ncu_d.log 2>&1
tail -3 gpurun_out/ncu_d.log; ls -la gpurun_out/*.ncu-rep
