# round 2, call q (1 GPU): the bench lines and ncu captures of call m again, with the captures exported to CSV on the box
# (the .ncu-rep files of call m pushed gpurun_out/ over the 64 MiB merge limit)
set -x; mkdir -p gpurun_out; T=r02q
timeout 600 python bench.py > gpurun_out/${T}_bench_cfg4_1gpu_default.json 2> gpurun_out/${T}_bench_cfg4.err
timeout 600 python bench.py --impl reference > gpurun_out/${T}_bench_reference_arm.json 2> /dev/null
timeout 900 python bench.py --config 2 > gpurun_out/${T}_bench_cfg2_1gpu.json 2> /dev/null
timeout 900 python bench.py --config 2 --impl reference > gpurun_out/${T}_bench_cfg2_reference_arm.json 2> /dev/null
timeout 900 python bench.py --config 3 --cpu-seconds 5 > gpurun_out/${T}_bench_cfg3_1gpu.json 2> /dev/null
timeout 900 python bench.py --config 5 --cpu-seconds 5 > gpurun_out/${T}_bench_cfg5_1gpu.json 2> /dev/null
timeout 900 python bench.py --config 4c --no-cpu-baseline > gpurun_out/${T}_bench_cfg4c_1gpu.json 2> /dev/null
timeout 600 python bench.py --math exact --no-cpu-baseline > gpurun_out/${T}_bench_cfg4_1gpu_exact.json 2> /dev/null
for c in 4 5 3; do
  timeout 400 python bench.py --config $c --steps 40 --warmup 3 --kernel-only --no-autotune > gpurun_out/${T}_plain_cfg$c.json 2>/dev/null || continue
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/${T}_launches_bench_cfg$c.csv python bench.py --config $c --steps 40 --warmup 3 --kernel-only --no-autotune > /dev/null 2>&1
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:step_rows2 -s 30 -c 2 -o /tmp/${T}_rows2_cfg$c python bench.py --config $c --steps 40 --warmup 3 --kernel-only --no-autotune > /dev/null 2>&1
  ncu -i /tmp/${T}_rows2_cfg$c.ncu-rep --page raw --csv > gpurun_out/${T}_step_rows2_cfg${c}_ncu_raw.csv 2>/dev/null
  ncu -i /tmp/${T}_rows2_cfg$c.ncu-rep --page details --csv > gpurun_out/${T}_step_rows2_cfg${c}_ncu_details.csv 2>/dev/null
  ncu -i /tmp/${T}_rows2_cfg$c.ncu-rep --page source --csv --print-source cuda,sass > /tmp/${T}_src_cfg$c.csv 2>/dev/null
  python scripts/ncu_lines.py /tmp/${T}_src_cfg$c.csv step_rows2 45 > gpurun_out/${T}_step_rows2_cfg${c}_source_hotspots.txt 2>&1
done
du -sh gpurun_out
python - <<'PY'
import json, glob
for f in sorted(glob.glob('gpurun_out/r02q_bench_*.json')):
    try:
        d = json.loads([l for l in open(f) if l.startswith('{')][0])
        r = d.get('roofline', {})
        print(f.split('r02q_bench_')[1], '%.3e' % d['value'], 'ms/step', d.get('ms_per_step'), 'e2e %.3e' % d['e2e']['value'], r.get('kernel', '')[:18], r.get('avg_launch_ms'), r.get('frac'))
    except Exception as e:
        print(f, 'FAILED', e)
PY
