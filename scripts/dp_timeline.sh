# per-launch timelines of two train() calls on 1 and 2 GPUs (march pipelined, launches serialised by the timing events)
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
A="--steps 128 --warmup 64 --no-cpu-baseline --no-ref-cuda"
python bench.py --gpus 1 $A --timeline gpurun_out/tl_n1 > /dev/null 2>&1
$TR bench.py --gpus 2 $A --timeline gpurun_out/tl_n2 > /dev/null 2>&1
ls gpurun_out/tl_*
