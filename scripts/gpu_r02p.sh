# round 2, call p (N GPUs = $1): cfg 4 strong scaling with the coalesced fused halo stores (+ slab suite when $2 = tests, + cfg 5 when $3 = cfg5)
set -x; mkdir -p gpurun_out; T=r02p; N=$1
B="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29561 bench.py --gpus $N --steps 200 --warmup 20"
timeout 900 $B > gpurun_out/${T}_bench_cfg4_${N}gpu.json 2> gpurun_out/${T}_bench_cfg4_${N}gpu.err
LGB200_HALO_FUSED=0 timeout 900 $B > gpurun_out/${T}_bench_cfg4_${N}gpu_push.json 2> gpurun_out/${T}_bench_cfg4_${N}gpu_push.err
if [ "$3" = "cfg5" ]; then timeout 1500 $B --config 5 > gpurun_out/${T}_bench_cfg5_${N}gpu.json 2> gpurun_out/${T}_bench_cfg5_${N}gpu.err; fi
if [ "$2" = "tests" ]; then timeout 1200 python -m pytest tests/test_slab_gpu.py -q --tb=line 2>&1 | grep -v "^\[W\|NCCL\|^frame" | tail -8 > gpurun_out/${T}_tests_slab_${N}gpu.log; cat gpurun_out/${T}_tests_slab_${N}gpu.log; fi
python - <<PY
import json, glob
for f in sorted(glob.glob("gpurun_out/r02p_bench_*_${N}gpu*.json")):
    try:
        d = json.loads([l for l in open(f) if l.startswith("{")][0])
        print(f.split("r02p_bench_")[1], "ms/step", d["ms_per_step"], "value %.3e" % d["value"], "e2e %.3e" % d["e2e"]["value"], "kernel ms", d["roofline"]["avg_launch_ms"], "share", d["roofline"]["share_of_step"])
    except Exception as e:
        print(f, "FAILED", e)
PY
