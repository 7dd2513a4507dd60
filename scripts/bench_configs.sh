#!/bin/bash
# The other BASELINE.json shapes (parity-test cases, not the bench line): render sweep of cfg1 / cfg4 / cfg5 and the
# cfg5 training step, both arms, one B200.  Writes gpurun_out/r02_cfg_*.json; round-2 copies are in profiles/.
mkdir -p gpurun_out
for w in cfg1 cfg4 cfg5; do
  for impl in ours reference; do
    python bench.py --workload $w --impl $impl --steps 3 --warmup 3 --no-cpu-baseline 2>gpurun_out/r02_cfg_${w}_${impl}.err | grep "^{" > gpurun_out/r02_cfg_${w}_${impl}.json
  done
done
for impl in ours reference; do
  python bench.py --mode train --workload cfg5 --impl $impl --steps 3 --warmup 3 2>gpurun_out/r02_cfg_cfg5train_${impl}.err | grep "^{" > gpurun_out/r02_cfg_cfg5train_${impl}.json
done
python bench.py --mode train --impl reference --steps 10 --warmup 3 2>/dev/null | grep "^{" > gpurun_out/r02_cfg_cfg3train_reference.json
python bench.py --impl reference --steps 10 --warmup 3 2>/dev/null | grep "^{" > gpurun_out/r02_cfg_cfg2_reference.json
