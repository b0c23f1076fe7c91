# A/B of one env switch on one box: usage  bash scripts/ab.sh VAR  (runs bench with VAR unset and VAR=1, twice each)
A="--steps 128 --warmup 96 --no-cpu-baseline --no-ref-cuda"
for rep in 1 2; do
python bench.py $A 2>/dev/null | grep '^{"metric' > gpurun_out/ab_new_$rep.json
env $1 python bench.py $A 2>/dev/null | grep '^{"metric' > gpurun_out/ab_old_$rep.json
done
for f in gpurun_out/ab_*.json; do python -c "
import json,sys
d=json.load(open('$f')); k=d['roofline']['kernels_ms_per_step']; print('$f', round(d['value']/1e6,1), round(d['ms_per_step'],4), {n: round(v*1000,1) for n,v in list(k.items())[:5]})"; done
