# round 2, call k (N GPUs = $1): strong scaling cfg 4, weak scaling cfg 5 (1.25 M particles per GPU), fused vs separate halo push
set -x; mkdir -p gpurun_out; T=r02k; N=$1
B="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $N --steps 200 --warmup 20"
timeout 900 $B > gpurun_out/${T}_bench_cfg4_${N}gpu.json 2> gpurun_out/${T}_bench_cfg4_${N}gpu.err
LGB200_HALO_FUSED=0 timeout 900 $B > gpurun_out/${T}_bench_cfg4_${N}gpu_push.json 2> gpurun_out/${T}_bench_cfg4_${N}gpu_push.err
timeout 1500 $B --config 5 > gpurun_out/${T}_bench_cfg5_${N}gpu.json 2> gpurun_out/${T}_bench_cfg5_${N}gpu.err
if [ "$2" = "tests" ]; then timeout 1200 python -m pytest tests/test_slab_gpu.py -q --tb=line 2>&1 | grep -v "^\[W\|NCCL\|^frame" | tail -8 > gpurun_out/${T}_tests_slab_${N}gpu.log; cat gpurun_out/${T}_tests_slab_${N}gpu.log; fi
python - <<PY
import json
for c in ("cfg4_${N}gpu", "cfg4_${N}gpu_push", "cfg5_${N}gpu"):
    try:
        d = json.loads([l for l in open("gpurun_out/r02k_bench_%s.json" % c) if l.startswith("{")][0])
        print(c, "ms/step", d["ms_per_step"], "value %.3e" % d["value"], "e2e %.3e" % d["e2e"]["value"], "kernel ms", d["roofline"]["avg_launch_ms"], "share", d["roofline"]["share_of_step"], d["config"]["particles"], d["config"].get("decomposition", "")[:110])
    except Exception as e:
        print(c, "FAILED", e)
PY
tail -3 gpurun_out/${T}_bench_cfg5_${N}gpu.err
