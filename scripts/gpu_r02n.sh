# round 2, call n (2 GPUs): relaxed-arithmetic contact law in the row kernels' marked-pair pass, coalesced fused halo stores
set -x; mkdir -p gpurun_out; T=r02n
timeout 1500 python -m pytest tests/test_slab_gpu.py tests/test_gpu_rows_kernel.py tests/test_gpu_fast_math.py tests/test_gpu_reorder.py -q --tb=line 2>&1 | grep -v "^\[W\|NCCL\|^frame" | tail -12 > gpurun_out/${T}_tests.log
: > gpurun_out/${T}_tune.jsonl
for c in 4 4c 5d 5; do timeout 400 python bench.py --config $c --steps 100 --warmup 10 --kernel-only >> gpurun_out/${T}_tune.jsonl 2>> gpurun_out/${T}_tune.err; done
B="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29551 bench.py --gpus 2 --steps 200 --warmup 20"
timeout 600 $B > gpurun_out/${T}_bench_cfg4_2gpu.json 2> gpurun_out/${T}_bench_cfg4_2gpu.err
python - <<'PY'
import json
for l in open('gpurun_out/r02n_tune.jsonl'):
    try: d=json.loads(l)
    except Exception: print(l[:200]); continue
    print(d.get('config'), 'shape', d.get('rows_shape'), 'C', d['contacts'], d['ms_per_step'], d['roofline']['kernel'][:18], d['roofline']['avg_launch_ms'], d['roofline']['frac'])
d = json.loads([l for l in open('gpurun_out/r02n_bench_cfg4_2gpu.json') if l.startswith('{')][0])
print('2gpu', d['ms_per_step'], '%.3e' % d['value'], 'e2e %.3e' % d['e2e']['value'], d['roofline']['avg_launch_ms'], d['roofline']['share_of_step'])
PY
cat gpurun_out/${T}_tests.log; tail -3 gpurun_out/${T}_tune.err
