# round 2, call i (2 GPUs): slab tests with the halo stores fused into step_rows2_kernel, device bond detection, late skin check;
# 2-GPU bench fused vs separate push kernel
set -x; mkdir -p gpurun_out; T=r02i
timeout 1200 python -m pytest tests/test_slab_gpu.py tests/test_gpu_reorder.py -q -x 2>&1 | tail -30 > gpurun_out/${T}_tests.log
B="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 2 --steps 200 --warmup 20"
timeout 600 $B > gpurun_out/${T}_bench_2gpu_fused.json 2> gpurun_out/${T}_bench_2gpu_fused.err
LGB200_HALO_FUSED=0 timeout 600 $B > gpurun_out/${T}_bench_2gpu_push.json 2> gpurun_out/${T}_bench_2gpu_push.err
timeout 300 python bench.py --steps 100 --warmup 10 --kernel-only > gpurun_out/${T}_cfg4_kernel.json 2>&1
tail -30 gpurun_out/${T}_tests.log; cut -c1-700 gpurun_out/${T}_bench_2gpu_fused.json; cut -c1-700 gpurun_out/${T}_bench_2gpu_push.json; tail -3 gpurun_out/${T}_bench_2gpu_fused.err; cut -c1-400 gpurun_out/${T}_cfg4_kernel.json
