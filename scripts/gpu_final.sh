# final check of the round on one GPU: full GPU test suite, smoke(), both bench arms
set -x; mkdir -p gpurun_out; T=r1z
python -m pytest tests -m gpu -x -q 2>&1 | tail -6 > gpurun_out/${T}_tests.log
python scripts/rebuild_ab.py > gpurun_out/${T}_rebuild_ab.jsonl 2>&1  # neighbor rows: current kernels vs the round-1 kernels
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${T}_smoke.log 2>&1
python bench.py --impl reference --steps 200 --warmup 20 > gpurun_out/${T}_bench_reference.json 2> gpurun_out/${T}_bench_reference.err
python bench.py > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err
cat gpurun_out/${T}_tests.log gpurun_out/${T}_smoke.log gpurun_out/${T}_bench_reference.json gpurun_out/${T}_bench.json
