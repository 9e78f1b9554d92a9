# timing of the default library and of every variant under libgran_b200/variants (scripts/build_variants.sh);
# FILLS="0.82 0.9" additionally sweeps the tile-fill knob on the default library
set -x; mkdir -p gpurun_out
: > gpurun_out/tune.jsonl
python bench.py --steps 100 --warmup 10 --kernel-only >> gpurun_out/tune.jsonl 2>> gpurun_out/tune.err
for f in $FILLS; do
  LGB200_TILE_FILL=$f LGB200_TAG=fill$f python bench.py --steps 100 --warmup 10 --kernel-only >> gpurun_out/tune.jsonl 2>> gpurun_out/tune.err
done
for v in libgran_b200/variants/*.so; do
  LGB200_LIB=$PWD/$v python bench.py --steps 100 --warmup 10 --kernel-only >> gpurun_out/tune.jsonl 2>> gpurun_out/tune.err
done
python - <<'PY'
import json
for l in open('gpurun_out/tune.jsonl'):
    d=json.loads(l); print(d['lib'].split('/')[-1], d.get('tag'), d['math'], d['ms_per_step'], d['roofline']['avg_launch_ms'])
PY
