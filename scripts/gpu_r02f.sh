# round 2, call f: section-8f variants on the GPU (alt_sinter, Hamaker variants, forward Euler, fixed particle, central attraction;
# header templates incl. the unmodified test/alt_sintered.cpp) + tile-shape variants of step_rows2_kernel
set -x; mkdir -p gpurun_out; T=r02f
timeout 900 python -m pytest tests/test_variants.py tests/test_cpp_host.py -m gpu -q 2>&1 | tail -30 > gpurun_out/${T}_tests_variants.log
: > gpurun_out/${T}_tune.jsonl
timeout 300 python bench.py --steps 100 --warmup 10 --kernel-only >> gpurun_out/${T}_tune.jsonl 2>> gpurun_out/${T}_tune.err
for v in libgran_b200/variants/*.so; do
  LGB200_LIB=$PWD/$v timeout 300 python bench.py --steps 100 --warmup 10 --kernel-only >> gpurun_out/${T}_tune.jsonl 2>> gpurun_out/${T}_tune.err
done
for c in 5 3 2; do LGB200_KERNEL_VARIANT=2 timeout 300 python bench.py --config $c --variant 2 --steps 100 --warmup 10 --kernel-only >> gpurun_out/${T}_tune.jsonl 2>> gpurun_out/${T}_tune.err; done
python - <<'PY'
import json
for l in open('gpurun_out/r02f_tune.jsonl'):
    try: d=json.loads(l)
    except Exception: print(l[:200]); continue
    print(d.get('config'), d['lib'].split('/')[-1], d['math'], d['ms_per_step'], d['roofline']['kernel'][:20], d['roofline']['avg_launch_ms'], d['roofline']['frac'])
PY
tail -30 gpurun_out/${T}_tests_variants.log; tail -5 gpurun_out/${T}_tune.err
