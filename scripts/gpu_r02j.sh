# round 2, call j (2 GPUs): slab suite (pinned tile shape, facets, fused halo), Newton-iteration variants of the relaxed math
set -x; mkdir -p gpurun_out; T=r02j
timeout 1500 python -m pytest tests/test_slab_gpu.py -q 2>&1 | tail -30 > gpurun_out/${T}_tests_slab.log
: > gpurun_out/${T}_tune.jsonl
timeout 300 python bench.py --steps 100 --warmup 10 --kernel-only >> gpurun_out/${T}_tune.jsonl 2>> gpurun_out/${T}_tune.err
for v in libgran_b200/variants/*.so; do
  LGB200_LIB=$PWD/$v timeout 300 python bench.py --steps 100 --warmup 10 --kernel-only >> gpurun_out/${T}_tune.jsonl 2>> gpurun_out/${T}_tune.err
done
LGB200_LIB=$PWD/libgran_b200/variants/liblgb200_nr1b2.so timeout 600 python -m pytest tests/test_gpu_rows_kernel.py tests/test_gpu_fast_math.py tests/test_gpu_reorder.py -q 2>&1 | tail -15 > gpurun_out/${T}_tests_nr1b2.log
B="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 2 --steps 200 --warmup 20"
timeout 600 $B > gpurun_out/${T}_bench_2gpu.json 2> gpurun_out/${T}_bench_2gpu.err
python - <<'PY'
import json
for l in open('gpurun_out/r02j_tune.jsonl'):
    try: d=json.loads(l)
    except Exception: print(l[:200]); continue
    print(d.get('config'), d['lib'].split('/')[-1], d['ms_per_step'], d['roofline']['avg_launch_ms'], d['roofline']['frac'])
PY
tail -30 gpurun_out/${T}_tests_slab.log; tail -15 gpurun_out/${T}_tests_nr1b2.log; cut -c1-300 gpurun_out/${T}_bench_2gpu.json; grep -o '"e2e": {[^}]*}' gpurun_out/${T}_bench_2gpu.json
