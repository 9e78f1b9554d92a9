# round 2, call t (1 GPU, short): A/B of the rebuilt neighbor-row kernels (rows identical to the round-1 kernels, ms per rebuild),
# their effect on the bench line of configs 5 and 4, then the oracle parity file with the new kernels as default
set -x; mkdir -p gpurun_out; T=r02t
timeout 130 python scripts/rebuild_ab.py > gpurun_out/${T}_rebuild_ab.jsonl 2> gpurun_out/${T}_rebuild_ab.err; echo "ab rc=$?" >> gpurun_out/${T}_rebuild_ab.jsonl
timeout 70 python bench.py --config 5 --steps 100 --warmup 10 --kernel-only > gpurun_out/${T}_bench_cfg5_v2.json 2> gpurun_out/${T}_bench_cfg5_v2.err
LGB200_REBUILD_V1=1 timeout 70 python bench.py --config 5 --steps 100 --warmup 10 --kernel-only > gpurun_out/${T}_bench_cfg5_v1.json 2> gpurun_out/${T}_bench_cfg5_v1.err
timeout 60 python bench.py --steps 100 --warmup 10 --kernel-only > gpurun_out/${T}_bench_cfg4_v2.json 2> gpurun_out/${T}_bench_cfg4_v2.err
timeout 120 python -m pytest tests/test_gpu_parity.py -q -m gpu -x 2>&1 | tail -n 8 > gpurun_out/${T}_tests_parity.log
cat gpurun_out/${T}_rebuild_ab.jsonl; tail -n 3 gpurun_out/${T}_rebuild_ab.err
for f in cfg5_v2 cfg5_v1 cfg4_v2; do python - "$f" <<'PY'
import json,sys
f=sys.argv[1]
try:
    d=json.loads(open(f'gpurun_out/r02t_bench_{f}.json').read().strip().splitlines()[-1])
    print(f, d['ms_per_step'], d['value'], d['roofline']['avg_launch_ms'])
except Exception as e: print(f, 'no line', e)
PY
done
cat gpurun_out/${T}_tests_parity.log
