# one data-parallel bench line at N GPUs: usage  bash scripts/dp_scale.sh N [extra bench args]
N=$1; shift
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 128 --warmup 64 --no-cpu-baseline --no-ref-cuda "$@" 2> gpurun_out/dp${N}.err | grep '^{"metric' > gpurun_out/bench_dp${N}.json
python -c "
import json
d=json.load(open('gpurun_out/bench_dp${N}.json')); k=d['roofline']['kernels_ms_per_step']; print($N, round(d['value']/1e6,1), round(d['ms_per_step'],4), d['samples_marched_per_step'], d['clocks'], {n: round(v*1000,1) for n,v in k.items() if n.startswith('peer') or n.startswith('adam')})"
