# usage: gpu_ncu_step.sh TAG  -- ncu --set full of two launches of the fused step kernel (fast mode bench)
set -x; mkdir -p gpurun_out; TAG=$1
python bench.py --steps 40 --warmup 3 --kernel-only > gpurun_out/${TAG}_plain.json 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:step_ -s 30 -c 2 -o gpurun_out/${TAG}_step python bench.py --steps 40 --warmup 3 --kernel-only > gpurun_out/${TAG}_ncu.log 2>&1
