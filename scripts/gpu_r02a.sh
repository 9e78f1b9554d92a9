# round 2, first GPU call: full GPU test suite (incl. the row-kernel-vs-oracle tests), smoke, one bench line per config,
# reference arms, launch list + full ncu capture of the row kernel (cfg 4) and of the tile kernel (cfg 5d)
set -x; mkdir -p gpurun_out; T=r02a
python -m pytest tests -m gpu -q 2>&1 | tail -40 > gpurun_out/${T}_tests.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${T}_smoke.log 2>&1
for c in 4 4c 2 3 5 5d; do
  python bench.py --config $c --cpu-seconds 4 > gpurun_out/${T}_bench_cfg$c.json 2> gpurun_out/${T}_bench_cfg$c.err
done
python bench.py --config 4 --math exact --no-cpu-baseline > gpurun_out/${T}_bench_cfg4_exact.json 2> gpurun_out/${T}_bench_cfg4_exact.err
python bench.py --config 5d --math exact --no-cpu-baseline > gpurun_out/${T}_bench_cfg5d_exact.json 2> gpurun_out/${T}_bench_cfg5d_exact.err
python bench.py --impl reference --config 2 > gpurun_out/${T}_ref_cfg2.json 2> gpurun_out/${T}_ref_cfg2.err
python bench.py --impl reference --config 4 > gpurun_out/${T}_ref_cfg4.json 2> gpurun_out/${T}_ref_cfg4.err
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/${T}_smi.csv
lscpu | head -20 > gpurun_out/${T}_lscpu.txt
python bench.py --steps 40 --warmup 3 --kernel-only > gpurun_out/${T}_plain4.json 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 200 -c 300 --csv --log-file gpurun_out/${T}_launches_cfg4.csv python bench.py --steps 40 --warmup 3 --kernel-only > gpurun_out/${T}_ncu_l4.log 2>&1
python bench.py --steps 40 --warmup 3 --kernel-only > gpurun_out/${T}_plain4b.json 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:step_rows -s 30 -c 2 -o gpurun_out/${T}_rows_cfg4 python bench.py --steps 40 --warmup 3 --kernel-only > gpurun_out/${T}_ncu_f4.log 2>&1
python bench.py --config 5d --steps 40 --warmup 3 --kernel-only > gpurun_out/${T}_plain5d.json 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:step_tiled -s 30 -c 2 -o gpurun_out/${T}_tiled_cfg5d python bench.py --config 5d --steps 40 --warmup 3 --kernel-only > gpurun_out/${T}_ncu_f5d.log 2>&1
tail -30 gpurun_out/${T}_tests.log; cat gpurun_out/${T}_smoke.log; cat gpurun_out/${T}_bench_cfg*.json | cut -c1-600
