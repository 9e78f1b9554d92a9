# usage: gpu_slab_check.sh N  -- slab-decomposition parity tests, then the torchrun bench on N GPUs with both halo transports
set -x; mkdir -p gpurun_out; N=$1
timeout 900 python -m pytest tests/test_slab_gpu.py -m gpu -x -q 2>&1 | tail -5 > gpurun_out/slab_tests_n$N.log
cat gpurun_out/slab_tests_n$N.log
for T in ${TRANSPORTS:-p2p nccl}; do
  LGB200_HALO=$T timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 200 --warmup 20 > gpurun_out/slab_bench_n${N}_$T.json 2> gpurun_out/slab_bench_n${N}_$T.err
  tail -3 gpurun_out/slab_bench_n${N}_$T.err; cat gpurun_out/slab_bench_n${N}_$T.json
done
