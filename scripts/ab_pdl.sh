#!/bin/bash
# A/B of programmatic dependent launch (csrc/common.cuh pdl_grid_sync / launch_pdl) on one B200.
#   FDLS_PDL=0                      plain stream-ordered launches
#   FDLS_PDL=1                      attribute + griddepcontrol.wait (the default build)
#   FDLS_PDL=1 + libfdls_early.so   additionally griddepcontrol.launch_dependents at the top of every kernel; build it with
#       FDLS_LIB=$PWD/4dlangsplat_b200/libfdls_early.so FDLS_NVCC_EXTRA="-DFDLS_PDL_EARLY=1" python 4dlangsplat_b200/build.py
# Prints: value, value_one_stream, e2e, stages_ms of the render sweep; then the cfg3 train object.  Results of round 2:
# profiles/r02_ab_pdl.txt.
for cfg in "FDLS_PDL=0" "FDLS_PDL=1" "FDLS_PDL=1 FDLS_LIB=$PWD/4dlangsplat_b200/libfdls_early.so"; do
  echo "== $cfg"
  env $cfg python bench.py --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print(d['value'], d['value_one_stream'], d['e2e']['value'], d['stages_ms'])
t=d['train']; print(t['value'], t['e2e']['value'], t['stages_ms'])
"
done
