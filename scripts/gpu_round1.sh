set -x; mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > gpurun_out/r1_tests.log
python bench.py --steps 200 --warmup 20 > gpurun_out/r1_bench.json 2> gpurun_out/r1_bench.err
python bench.py --steps 200 --warmup 20 --variant 0 --no-cpu-baseline > gpurun_out/r1_bench_v0.json 2>> gpurun_out/r1_bench.err
python bench.py --impl reference --steps 20 --warmup 2 > gpurun_out/r1_bench_ref.json 2>> gpurun_out/r1_bench.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1_launches.csv python bench.py --steps 40 --warmup 3 --no-cpu-baseline > gpurun_out/r1_ncu.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:step_tiled_kernel -s 30 -c 2 -o gpurun_out/r1_step_tiled python bench.py --steps 40 --warmup 3 --no-cpu-baseline > gpurun_out/r1_ncu_full.log 2>&1
nvidia-smi > gpurun_out/r1_smi.txt; nproc >> gpurun_out/r1_smi.txt
