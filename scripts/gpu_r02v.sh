# round 2, call v (1 GPU, short): row sweeps reading a cell-ordered copy of x (default) against the same kernels without it
# (LGB200_REBUILD_XS=0) and the round-1 kernels (LGB200_REBUILD_V1=1): rows identical, bench lines of configs 5, 3, 4
set -x; mkdir -p gpurun_out; T=r02v
timeout 100 python scripts/rebuild_ab.py > gpurun_out/${T}_rebuild_ab.jsonl 2> gpurun_out/${T}_rebuild_ab.err; echo "ab rc=$?" >> gpurun_out/${T}_rebuild_ab.jsonl
B="python bench.py --steps 100 --warmup 10 --kernel-only"
timeout 40 $B --config 5 > gpurun_out/${T}_bench_cfg5_xs.json 2> gpurun_out/${T}_bench_cfg5_xs.err
LGB200_REBUILD_XS=0 timeout 40 $B --config 5 > gpurun_out/${T}_bench_cfg5_noxs.json 2> gpurun_out/${T}_bench_cfg5_noxs.err
LGB200_LIB=$PWD/libgran_b200/variants/liblgb200_b8.so timeout 40 $B --config 5 > gpurun_out/${T}_bench_cfg5_xsb8.json 2> gpurun_out/${T}_bench_cfg5_xsb8.err
LGB200_LIB=$PWD/libgran_b200/variants/liblgb200_b2.so timeout 40 $B --config 5 > gpurun_out/${T}_bench_cfg5_xsb2.json 2> gpurun_out/${T}_bench_cfg5_xsb2.err
timeout 40 $B --config 3 > gpurun_out/${T}_bench_cfg3_xs.json 2> gpurun_out/${T}_bench_cfg3_xs.err
LGB200_REBUILD_V1=1 timeout 40 $B --config 3 > gpurun_out/${T}_bench_cfg3_v1.json 2> gpurun_out/${T}_bench_cfg3_v1.err
timeout 40 $B > gpurun_out/${T}_bench_cfg4_xs.json 2> gpurun_out/${T}_bench_cfg4_xs.err
LGB200_REBUILD_V1=1 timeout 40 $B > gpurun_out/${T}_bench_cfg4_v1.json 2> gpurun_out/${T}_bench_cfg4_v1.err
timeout 60 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'row_sweep|retag|tag_clear|gather3|cell_' --csv --log-file gpurun_out/${T}_launches_rebuild_cfg5.csv python bench.py --config 5 --steps 40 --warmup 10 --kernel-only > /dev/null 2> gpurun_out/${T}_ncu.err
cat gpurun_out/${T}_rebuild_ab.jsonl; tail -n 3 gpurun_out/${T}_rebuild_ab.err
python - <<'PY'
import json, glob, csv, collections
for f in sorted(glob.glob('gpurun_out/r02v_bench_*.json')):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        print(f.split('bench_')[1], d['ms_per_step'], '%.4g' % d['value'], d['roofline']['avg_launch_ms'])
    except Exception as e:
        print(f, 'no line', e)
rows = list(csv.reader(open('gpurun_out/r02v_launches_rebuild_cfg5.csv'))); hdr = None; agg = collections.OrderedDict()
for r in rows:
    if 'Kernel Name' in r: hdr = {h: i for i, h in enumerate(r)}; continue
    if hdr is None or len(r) < len(hdr): continue
    nm = r[hdr['Kernel Name']].split('(')[0][-40:]; v = float(r[hdr['Metric Value']].replace(',', '')); u = r[hdr['Metric Unit']]
    agg.setdefault(nm, []).append(round(v / 1000 if u == 'ns' else v * 1000 if u == 'ms' else v, 1))
for nm, vs in agg.items(): print('   ', nm, vs[-4:])
PY
