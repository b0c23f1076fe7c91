# ncu evidence for one build: plain run first, then the launch list, then one --set full capture of the step's main kernels.
# usage: [WARM=128 SKIP=1150 COUNT=16] bash scripts/ncu_capture.sh TAG   (9 matching launches per step; the defaults land in early training)
TAG=${1:-v6}
B="python bench.py --steps 32 --warmup ${WARM:-48} --no-cpu-baseline --no-ref-cuda --no-extras"
$B > gpurun_out/ncu_plain_$TAG.json 2> gpurun_out/ncu_plain_$TAG.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -s 600 -c 500 --csv --log-file gpurun_out/launches_$TAG.csv $B > gpurun_out/ncu_ll_$TAG.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"hash_encode_fwd|hash_encode_bwd|mlp_kernel|mlp_fwd_kernel|mlp_train_kernel|adam_ema|composite_grad|composite_count|wgrad_reduce|march_count" -s ${SKIP:-420} -c ${COUNT:-18} -o gpurun_out/prof_$TAG -f $B > gpurun_out/ncu_full_$TAG.log 2>&1
ls -la gpurun_out/prof_$TAG.ncu-rep gpurun_out/launches_$TAG.csv
