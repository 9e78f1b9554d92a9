# round 2, call h: full GPU suite with the 8f variants, renumbering, dump pipeline; warp-cooperative turnaround + autotune timings
set -x; mkdir -p gpurun_out; T=r02h
timeout 1800 python -m pytest tests -m gpu -q -x --deselect "tests/test_cpp_host.py::test_reference_driver_reaches_its_targets_on_gpu[alt_sintered]" 2>&1 | tail -30 > gpurun_out/${T}_gpu_tests.log
: > gpurun_out/${T}_tune.jsonl
run() { timeout 400 python bench.py "$@" --steps 100 --warmup 10 --kernel-only >> gpurun_out/${T}_tune.jsonl 2>> gpurun_out/${T}_tune.err; }
run --config 4
LGB200_TAG=noautotune run --config 4 --no-autotune
for c in 4c 2 3 5 5d; do run --config $c; done
python - <<'PY'
import json
for l in open('gpurun_out/r02h_tune.jsonl'):
    try: d=json.loads(l)
    except Exception: print(l[:200]); continue
    print(d.get('config'), d.get('tag'), 'reorder', d.get('reorder'), 'shape', d.get('rows_shape'), 'C', d['contacts'], 'Bd', d['bonds_ordered'], d['ms_per_step'], d['roofline']['kernel'][:18], d['roofline']['avg_launch_ms'], d['roofline']['frac'])
PY
tail -30 gpurun_out/${T}_gpu_tests.log; tail -5 gpurun_out/${T}_tune.err
