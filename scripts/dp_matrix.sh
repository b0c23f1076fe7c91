# 2-GPU A/B on one box: usage  bash scripts/dp_matrix.sh VAR=VALUE  (default build vs the setting, twice each) + one 1-GPU line
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
A="--steps 128 --warmup 64 --no-cpu-baseline --no-ref-cuda"
python bench.py --gpus 1 $A 2>/dev/null | grep '^{"metric' > gpurun_out/dp_n1.json
for rep in 1 2; do
$TR bench.py --gpus 2 $A 2>/dev/null | grep '^{"metric' > gpurun_out/dp_n2_new_$rep.json
env $1 $TR bench.py --gpus 2 $A 2>/dev/null | grep '^{"metric' > gpurun_out/dp_n2_old_$rep.json
done
for f in gpurun_out/dp_n*.json; do python -c "
import json,sys
d=json.load(open('$f')); k=d['roofline']['kernels_ms_per_step']; print('$f', round(d['value']/1e6,1), round(d['ms_per_step'],4), d['samples_marched_per_step'], {n: round(v*1000,1) for n,v in k.items() if n.startswith('peer') or n.startswith('adam')})"; done
