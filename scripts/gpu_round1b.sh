set -x; mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > gpurun_out/r1b_tests.log
python bench.py --steps 200 --warmup 20 > gpurun_out/r1b_bench.json 2> gpurun_out/r1b_bench.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1b_launches.csv python bench.py --steps 40 --warmup 3 --no-cpu-baseline > gpurun_out/r1b_ncu.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:step_tiled_kernel -s 30 -c 2 -o gpurun_out/r1b_step_tiled python bench.py --steps 40 --warmup 3 --no-cpu-baseline > gpurun_out/r1b_ncu_full.log 2>&1
nvidia-smi > gpurun_out/r1b_smi.txt; nproc >> gpurun_out/r1b_smi.txt
