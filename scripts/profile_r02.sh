#!/bin/bash
# Round-2 profiles of the default bench command (cfg2 render sweep + the cfg3 train object), one B200:
#   launch list (ncu --metrics gpu__time_duration.sum) and one --set full capture of every kernel of a frame / a step.
# Each ncu run follows a plain run of the same command that exited 0.  Summaries: profiles/r02_launches_*.csv,
# profiles/r02_ncu_full_*.json (python profiles/ncu_summary.py).
export FDLS_BENCH_FRAMES_PER_STEP=2
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
mkdir -p gpurun_out
$CMD > gpurun_out/plain_r02.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -s 60 -c 400 --csv --log-file gpurun_out/r02_launches_ours_cfg2.csv $CMD > gpurun_out/ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"preprocess_fwd|rs_onesweep|scan_|part_|render_fwd" -s 20 -c 10 -o gpurun_out/r02_frame $CMD > gpurun_out/ncu_f.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"render_bwd|preprocess_bwd|sh_grad_expand" -s 4 -c 3 -o gpurun_out/r02_bwd $CMD > gpurun_out/ncu_b.log 2>&1
$CMD --impl reference > gpurun_out/plain_r02_ref.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -s 60 -c 200 --csv --log-file gpurun_out/r02_launches_reference_cfg2.csv $CMD --impl reference > gpurun_out/ncu_lr.log 2>&1
ls -la gpurun_out
