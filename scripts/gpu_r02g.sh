# round 2, call g: spatial renumbering + shape-templated step_rows2_kernel (<128,2>, <64,4>, <32,8>), lane-parallel marked pairs
set -x; mkdir -p gpurun_out; T=r02g
timeout 900 python -m pytest tests/test_gpu_reorder.py tests/test_gpu_rows_kernel.py tests/test_gpu_fast_math.py -q 2>&1 | tail -30 > gpurun_out/${T}_tests.log
: > gpurun_out/${T}_tune.jsonl
run() { timeout 400 python bench.py "$@" --steps 100 --warmup 10 --kernel-only >> gpurun_out/${T}_tune.jsonl 2>> gpurun_out/${T}_tune.err; }
run --config 4
for c in 5 3 2 5d 4c; do
  LGB200_TAG=rows2 run --config $c
  LGB200_TAG=rows2-noreorder run --config $c --reorder off
  LGB200_TAG=tile LGB200_ROWS_KMAX=24 run --config $c
done
for sh in 0 1 2; do LGB200_TAG=shape$sh LGB200_ROWS_SHAPE=$sh run --config 5; LGB200_TAG=shape$sh LGB200_ROWS_SHAPE=$sh run --config 3; LGB200_TAG=shape$sh LGB200_ROWS_SHAPE=$sh run --config 5d; done
python - <<'PY'
import json
for l in open('gpurun_out/r02g_tune.jsonl'):
    try: d=json.loads(l)
    except Exception: print(l[:200]); continue
    print(d.get('config'), d.get('tag'), 'reorder', d.get('reorder'), 'C', d['contacts'], d['ms_per_step'], d['roofline']['kernel'][:18], d['roofline']['avg_launch_ms'], d['roofline']['frac'])
PY
tail -30 gpurun_out/${T}_tests.log; tail -5 gpurun_out/${T}_tune.err
