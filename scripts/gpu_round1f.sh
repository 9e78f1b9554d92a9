# round-1f: final measurements of the round (row kernel for fast math, tile kernel for exact math)
set -x; mkdir -p gpurun_out; T=r1f
python -m pytest tests -m gpu -x -q 2>&1 | tail -8 > gpurun_out/${T}_tests.log
python bench.py --steps 200 --warmup 20 > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err
python bench.py --steps 200 --warmup 20 --math exact --no-cpu-baseline > gpurun_out/${T}_bench_exact.json 2> gpurun_out/${T}_bench_exact.err
python bench.py --impl reference --steps 200 --warmup 20 > gpurun_out/${T}_bench_reference.json 2> gpurun_out/${T}_bench_reference.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${T}_launches.csv python bench.py --steps 40 --warmup 3 --no-cpu-baseline > gpurun_out/${T}_ncu.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:step_rows_kernel -s 30 -c 2 -o gpurun_out/${T}_step_rows python bench.py --steps 40 --warmup 3 --kernel-only > gpurun_out/${T}_ncu_full.log 2>&1
nvidia-smi > gpurun_out/${T}_smi.txt; nproc >> gpurun_out/${T}_smi.txt
cat gpurun_out/${T}_tests.log gpurun_out/${T}_bench.json gpurun_out/${T}_bench_exact.json gpurun_out/${T}_bench_reference.json
