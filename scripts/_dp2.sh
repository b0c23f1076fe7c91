#!/bin/bash
for mode in peer nccl peer; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 128 --warmup 64 --dp-mode $mode 2>gpurun_out/bench2_$mode.err | tail -1 > gpurun_out/bench2_$mode.json
python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/bench2_$mode.json").read())
    print("$mode", round(d["value"]/1e6,1), round(d["ms_per_step"],4), d["loss"], d["config"]["parallelism"])
except Exception as e:
    print("$mode failed", e); print(open("gpurun_out/bench2_$mode.err").read()[-3000:])
PY
done
