#!/bin/bash
# usage: _sweep.sh ENVVAR v1 v2 ...
var=$1; shift
for v in "$@"; do
  export $var=$v
  timeout 300 python bench.py --steps 128 --warmup 64 --no-cpu-baseline --no-ref-cuda 2>/dev/null > /tmp/sweep.json
  python - <<PY
import json
d=json.loads(open("/tmp/sweep.json").read().strip().splitlines()[-1]); k=d["roofline"]["kernels_ms_per_step"]
print("$var=$v", round(d["value"]/1e6,1), round(d["ms_per_step"],4), round(sum(k.values()),4), {a:round(b*1000,1) for a,b in list(k.items())[:7]})
PY
done
