# correctness subset + timing of the default library and all variants
set -x; mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py tests/test_gpu_fast_math.py -m gpu -x -q 2>&1 | tail -3
bash scripts/gpu_tune.sh 2>/dev/null | grep -v "^+"
python bench.py --steps 100 --warmup 10 --kernel-only --math exact 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('exact', d['ms_per_step'], d['roofline']['avg_launch_ms'])"
