# round 2, call w (1 GPU, short): neighbor rebuild with a hit list between the count and fill passes (no second sweep):
# rows identical to the round-1 kernels, oracle parity file, bench lines of configs 5, 4, 3, per-kernel durations
set -x; mkdir -p gpurun_out; T=r02w
timeout 45 python scripts/rebuild_ab.py > gpurun_out/${T}_rebuild_ab.jsonl 2> gpurun_out/${T}_rebuild_ab.err; echo "ab rc=$?" >> gpurun_out/${T}_rebuild_ab.jsonl
timeout 30 python -m pytest tests/test_gpu_parity.py -q -m gpu -x 2>&1 | tail -n 6 > gpurun_out/${T}_tests_parity.log
B="python bench.py --steps 100 --warmup 10 --kernel-only"
timeout 25 $B --config 5 > gpurun_out/${T}_bench_cfg5.json 2> gpurun_out/${T}_bench_cfg5.err
timeout 25 $B > gpurun_out/${T}_bench_cfg4.json 2> gpurun_out/${T}_bench_cfg4.err
timeout 25 $B --config 3 > gpurun_out/${T}_bench_cfg3.json 2> gpurun_out/${T}_bench_cfg3.err
cat gpurun_out/${T}_rebuild_ab.jsonl; tail -n 3 gpurun_out/${T}_rebuild_ab.err; cat gpurun_out/${T}_tests_parity.log
python - <<'PY'
import json, glob
for f in sorted(glob.glob('gpurun_out/r02w_bench_*.json')):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        print(f.split('bench_')[1], d['ms_per_step'], '%.4g' % d['value'], d['roofline']['avg_launch_ms'])
    except Exception as e:
        print(f, 'no line', e)
PY
timeout 30 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'row_|retag|tag_clear' --csv --log-file gpurun_out/${T}_launches_rebuild_cfg5.csv python bench.py --config 5 --steps 40 --warmup 10 --kernel-only > /dev/null 2> gpurun_out/${T}_ncu.err
grep -E "row_|retag|tag_clear" gpurun_out/${T}_launches_rebuild_cfg5.csv | tail -n 8 | awk -F'","' '{print $5, $NF}'
