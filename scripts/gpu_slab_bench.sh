# usage: gpu_slab_bench.sh N [copies]   -- torchrun bench on N GPUs (p2p halo transport)
set -x; mkdir -p gpurun_out
N=$1; COPIES=${2:-6667}
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 200 --warmup 20 --copies $COPIES > gpurun_out/slab_bench_n$N.json 2> gpurun_out/slab_bench_n$N.err
tail -5 gpurun_out/slab_bench_n$N.err; cat gpurun_out/slab_bench_n$N.json
