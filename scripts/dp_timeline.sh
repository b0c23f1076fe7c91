# per-launch timelines of two train() calls at N GPUs (march pipelined, launches serialised by the timing events): bash scripts/dp_timeline.sh N
N=${1:-2}
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 128 --warmup 64 --no-cpu-baseline --no-ref-cuda --timeline gpurun_out/tl_n$N 2> gpurun_out/tl.err | grep '^{"metric' > gpurun_out/tl_bench_n$N.json
ls gpurun_out/tl_n$N*
