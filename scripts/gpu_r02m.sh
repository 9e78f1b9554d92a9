# round 2, call m (1 GPU): final state -- full GPU suite, smoke, the default bench line + reference arm, full bench lines of
# configs 2 / 3 / 5 / 4c, launch list and --set full captures of the step kernel on configs 4, 5, 3, 2
set -x; mkdir -p gpurun_out; T=r02m
timeout 1800 python -m pytest tests -m gpu -q 2>&1 | tail -12 > gpurun_out/${T}_gpu_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${T}_smoke.log 2>&1
timeout 600 python bench.py > gpurun_out/${T}_bench_cfg4_1gpu_default.json 2> gpurun_out/${T}_bench_cfg4.err
timeout 600 python bench.py --impl reference > gpurun_out/${T}_bench_reference_arm.json 2> gpurun_out/${T}_bench_reference_arm.err
timeout 900 python bench.py --config 2 > gpurun_out/${T}_bench_cfg2_1gpu.json 2> gpurun_out/${T}_bench_cfg2.err
timeout 900 python bench.py --config 2 --impl reference > gpurun_out/${T}_bench_cfg2_reference_arm.json 2> gpurun_out/${T}_bench_cfg2_reference_arm.err
timeout 900 python bench.py --config 3 --cpu-seconds 5 > gpurun_out/${T}_bench_cfg3_1gpu.json 2> gpurun_out/${T}_bench_cfg3.err
timeout 900 python bench.py --config 5 --cpu-seconds 5 > gpurun_out/${T}_bench_cfg5_1gpu.json 2> gpurun_out/${T}_bench_cfg5.err
timeout 900 python bench.py --config 4c --no-cpu-baseline > gpurun_out/${T}_bench_cfg4c_1gpu.json 2> gpurun_out/${T}_bench_cfg4c.err
timeout 600 python bench.py --math exact --no-cpu-baseline > gpurun_out/${T}_bench_cfg4_1gpu_exact.json 2> gpurun_out/${T}_bench_cfg4_exact.err
# profiles: only after the same command has exited 0 without ncu
for c in 4 5 3 2; do
  timeout 400 python bench.py --config $c --steps 40 --warmup 3 --kernel-only --no-autotune > gpurun_out/${T}_plain_cfg$c.json 2>&1 || continue
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/${T}_launches_bench_cfg$c.csv python bench.py --config $c --steps 40 --warmup 3 --kernel-only --no-autotune > gpurun_out/${T}_ncu_launches_cfg$c.log 2>&1
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:step_rows2 -s 30 -c 2 -o gpurun_out/${T}_rows2_cfg$c python bench.py --config $c --steps 40 --warmup 3 --kernel-only --no-autotune > gpurun_out/${T}_ncu_full_cfg$c.log 2>&1
done
tail -12 gpurun_out/${T}_gpu_tests.log; cat gpurun_out/${T}_smoke.log
python - <<'PY'
import json, glob
for f in sorted(glob.glob('gpurun_out/r02m_bench_*.json')):
    try:
        d = json.loads([l for l in open(f) if l.startswith('{')][0])
        r = d.get('roofline', {})
        print(f.split('r02m_bench_')[1], '%.3e' % d['value'], 'ms/step', d.get('ms_per_step'), 'e2e %.3e' % d['e2e']['value'], r.get('kernel', '')[:18], r.get('avg_launch_ms'), r.get('frac'), 'cpu', d.get('cpu_baseline', {}).get('value'), d.get('cpu_baseline', {}).get('kind'))
    except Exception as e:
        print(f, 'FAILED', e)
PY
