# round 2, call d: step_rows2_kernel with paired bridge evaluation + Newton square roots: parity, timing, ncu
set -x; mkdir -p gpurun_out; T=r02d
timeout 900 python -m pytest tests/test_gpu_rows_kernel.py tests/test_gpu_fast_math.py -q 2>&1 | tail -25 > gpurun_out/${T}_tests_rows2.log
: > gpurun_out/${T}_tune.jsonl
timeout 300 python bench.py --steps 100 --warmup 10 --kernel-only >> gpurun_out/${T}_tune.jsonl 2>> gpurun_out/${T}_tune.err
LGB200_ROWS_KERNEL=1 LGB200_TAG=rows1 timeout 300 python bench.py --steps 100 --warmup 10 --kernel-only >> gpurun_out/${T}_tune.jsonl 2>> gpurun_out/${T}_tune.err
for v in libgran_b200/variants/*.so; do
  LGB200_LIB=$PWD/$v timeout 300 python bench.py --steps 100 --warmup 10 --kernel-only >> gpurun_out/${T}_tune.jsonl 2>> gpurun_out/${T}_tune.err
done
for c in 2 3 5; do timeout 300 python bench.py --config $c --steps 100 --warmup 10 --kernel-only >> gpurun_out/${T}_tune.jsonl 2>> gpurun_out/${T}_tune.err; done
python - <<'PY'
import json
for l in open('gpurun_out/r02d_tune.jsonl'):
    try: d=json.loads(l)
    except Exception: print(l[:200]); continue
    print(d.get('config'), d['lib'].split('/')[-1], d.get('tag'), d['math'], d['ms_per_step'], d['roofline']['kernel'][:20], d['roofline']['avg_launch_ms'], d['roofline']['frac'])
PY
timeout 300 python bench.py --steps 40 --warmup 3 --kernel-only > gpurun_out/${T}_plain4.json 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:step_rows2 -s 30 -c 2 -o gpurun_out/${T}_rows2_cfg4 python bench.py --steps 40 --warmup 3 --kernel-only > gpurun_out/${T}_ncu_f4.log 2>&1
tail -25 gpurun_out/${T}_tests_rows2.log
