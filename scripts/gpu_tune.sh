set -x; mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -3 > gpurun_out/tune_tests.log
: > gpurun_out/tune.jsonl
python bench.py --steps 200 --warmup 20 --kernel-only >> gpurun_out/tune.jsonl 2>> gpurun_out/tune.err
for v in libgran_b200/variants/*.so; do
  LGB200_LIB=$PWD/$v python bench.py --steps 200 --warmup 20 --kernel-only >> gpurun_out/tune.jsonl 2>> gpurun_out/tune.err
done
