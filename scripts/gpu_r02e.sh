# round 2, call e (first call after the container was re-created): full GPU suite, smoke, default bench line, reference arm,
# kernel-only lines for every config, launch list and --set full capture of step_rows2_kernel
set -x; mkdir -p gpurun_out; T=r02e
timeout 1500 python -m pytest tests -m gpu -q 2>&1 | tail -30 > gpurun_out/${T}_gpu_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${T}_smoke.log 2>&1
timeout 600 python bench.py > gpurun_out/${T}_bench_cfg4_default.json 2> gpurun_out/${T}_bench_cfg4_default.err
timeout 600 python bench.py --impl reference > gpurun_out/${T}_bench_reference_arm.json 2> gpurun_out/${T}_bench_reference_arm.err
: > gpurun_out/${T}_tune.jsonl
LGB200_ROWS_KERNEL=1 LGB200_TAG=rows1 timeout 300 python bench.py --steps 100 --warmup 10 --kernel-only >> gpurun_out/${T}_tune.jsonl 2>> gpurun_out/${T}_tune.err
for c in 4 4c 2 3 5 5d; do timeout 400 python bench.py --config $c --steps 100 --warmup 10 --kernel-only >> gpurun_out/${T}_tune.jsonl 2>> gpurun_out/${T}_tune.err; done
for c in 4 5; do timeout 400 python bench.py --config $c --math exact --steps 100 --warmup 10 --kernel-only >> gpurun_out/${T}_tune.jsonl 2>> gpurun_out/${T}_tune.err; done
python - <<'PY'
import json
for l in open('gpurun_out/r02e_tune.jsonl'):
    try: d=json.loads(l)
    except Exception: print(l[:200]); continue
    print(d.get('config'), d.get('tag'), d['math'], d['particles'], d['pairs'], d['contacts'], d['ms_per_step'], d['roofline']['kernel'][:20], d['roofline']['avg_launch_ms'], d['roofline']['frac'])
PY
timeout 300 python bench.py --steps 40 --warmup 3 --kernel-only > gpurun_out/${T}_plain4.json 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${T}_launches_bench_cfg4.csv python bench.py --steps 40 --warmup 3 --kernel-only > gpurun_out/${T}_ncu_launches.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:step_rows2 -s 30 -c 2 -o gpurun_out/${T}_rows2_cfg4 python bench.py --steps 40 --warmup 3 --kernel-only > gpurun_out/${T}_ncu_f4.log 2>&1
tail -25 gpurun_out/${T}_gpu_tests.log; cat gpurun_out/${T}_smoke.log; cut -c1-600 gpurun_out/${T}_bench_cfg4_default.json; cut -c1-400 gpurun_out/${T}_bench_reference_arm.json; tail -5 gpurun_out/${T}_tune.err
