#!/bin/bash
# A/B of the default bench line: each argument is "label:ENV=VALUE[,ENV=VALUE...]" (or "label:" for the defaults)
mkdir -p gpurun_out
for spec in "$@"; do
  label=${spec%%:*}; envs=${spec#*:}
  ( IFS=,; for kv in $envs; do [ -n "$kv" ] && export "$kv"; done
    timeout 400 python bench.py --no-cpu-baseline > gpurun_out/ab_$label.json 2> gpurun_out/ab_$label.err )
  python - "$label" <<PY
import json, sys
d = json.loads(open("gpurun_out/ab_%s.json" % sys.argv[1]).read()); t = d.get("train", {})
print(sys.argv[1], "render", d["value"], "one_stream", d.get("value_one_stream"), "e2e", d["e2e"]["value"], "train", t.get("value"), "host_issue_ms", t.get("host_issue_ms_per_step"), t.get("stages_ms"))
PY
done
