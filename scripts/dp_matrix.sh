TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
A="--steps 128 --warmup 64 --no-cpu-baseline --no-ref-cuda"
for rep in 1 2 3; do
NGP_NO_CLOCKS=1 $TR bench.py --gpus 2 $A 2>/dev/null | grep '^{"metric' > gpurun_out/w_n2_noclk_$rep.json
$TR bench.py --gpus 2 $A 2>/dev/null | grep '^{"metric' > gpurun_out/w_n2_r0_$rep.json
NGP_CLOCKS_ALL_RANKS=1 $TR bench.py --gpus 2 $A 2>/dev/null | grep '^{"metric' > gpurun_out/w_n2_all_$rep.json
done
for f in gpurun_out/w_*.json; do python -c "
import json,sys
d=json.load(open('$f')); print('$f', round(d['value']/1e6,1), round(d['ms_per_step'],4), round(d['e2e']['wall_ms_per_step'],4), d['samples_marched_per_step'], d['clocks'])"; done
