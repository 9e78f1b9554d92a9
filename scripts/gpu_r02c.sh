# round 2, call c: single-stage step_rows2_kernel (2 CTAs/SM, L2 prefetch) + rebuild without host round trips: full GPU suite,
# timing against step_rows_kernel, tile-shape variants, cfg 4 / 4c lines, ncu capture
set -x; mkdir -p gpurun_out; T=r02c
timeout 1500 python -m pytest tests -m gpu -q 2>&1 | tail -25 > gpurun_out/${T}_tests_rows2.log
echo skipped > gpurun_out/${T}_tests_rows1.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${T}_smoke.log 2>&1
: > gpurun_out/${T}_tune.jsonl
timeout 300 python bench.py --steps 100 --warmup 10 --kernel-only >> gpurun_out/${T}_tune.jsonl 2>> gpurun_out/${T}_tune.err
LGB200_ROWS_KERNEL=1 LGB200_TAG=rows1 timeout 300 python bench.py --steps 100 --warmup 10 --kernel-only >> gpurun_out/${T}_tune.jsonl 2>> gpurun_out/${T}_tune.err
for v in libgran_b200/variants/*.so; do
  LGB200_LIB=$PWD/$v timeout 300 python bench.py --steps 100 --warmup 10 --kernel-only >> gpurun_out/${T}_tune.jsonl 2>> gpurun_out/${T}_tune.err
done
timeout 600 python bench.py --config 4c --cpu-seconds 4 > gpurun_out/${T}_bench_cfg4c.json 2> gpurun_out/${T}_bench_cfg4c.err
timeout 600 python bench.py --config 4 --cpu-seconds 4 > gpurun_out/${T}_bench_cfg4.json 2> gpurun_out/${T}_bench_cfg4.err
python - <<'PY'
import json
for l in open('gpurun_out/r02c_tune.jsonl'):
    try: d=json.loads(l)
    except Exception: print(l[:200]); continue
    print(d['lib'].split('/')[-1], d.get('tag'), d['math'], d['ms_per_step'], d['roofline']['kernel'][:20], d['roofline']['avg_launch_ms'], d['roofline']['frac'])
PY
timeout 300 python bench.py --steps 40 --warmup 3 --kernel-only > gpurun_out/${T}_plain4.json 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:step_rows2 -s 30 -c 2 -o gpurun_out/${T}_rows2_cfg4 python bench.py --steps 40 --warmup 3 --kernel-only > gpurun_out/${T}_ncu_f4.log 2>&1
tail -25 gpurun_out/${T}_tests_rows2.log; tail -8 gpurun_out/${T}_tests_rows1.log; cat gpurun_out/${T}_smoke.log; cut -c1-400 gpurun_out/${T}_bench_cfg4c.json
