# round 2, call u (1 GPU, short): per-kernel durations of the neighbor rebuild (ncu launch lists) -- new kernels, round-1 kernels,
# and four tuning builds of the new fill pass (no sort / 1 or 8 candidates in flight / 48-entry staging at 8 CTAs per SM)
set -x; mkdir -p gpurun_out; T=r02u
NCU="ncu --metrics gpu__time_duration.sum --clock-control none -k regex:row_sweep|retag|tag_clear --csv"
export AB_REPS=1
AB_CASES=packing_1.25M_rv5r,cfg4_1M_aggregates timeout 60 $NCU --log-file gpurun_out/${T}_launches_v2.csv python scripts/rebuild_ab.py child > /dev/null 2> gpurun_out/${T}_v2.err
AB_CASES=packing_1.25M_rv5r,cfg4_1M_aggregates LGB200_REBUILD_V1=1 timeout 60 $NCU --log-file gpurun_out/${T}_launches_v1.csv python scripts/rebuild_ab.py child > /dev/null 2> gpurun_out/${T}_v1.err
for v in nosort b1 b8 cap48; do
  AB_CASES=packing_1.25M_rv5r LGB200_LIB=$PWD/libgran_b200/variants/liblgb200_$v.so timeout 60 $NCU --log-file gpurun_out/${T}_launches_$v.csv python scripts/rebuild_ab.py child > /dev/null 2> gpurun_out/${T}_$v.err
done
python - <<'PY'
import csv, glob, collections
for f in sorted(glob.glob('gpurun_out/r02u_launches_*.csv')):
    rows = list(csv.reader(open(f)))
    hdr = None
    agg = collections.OrderedDict()
    for r in rows:
        if 'Kernel Name' in r: hdr = {h: i for i, h in enumerate(r)}; continue
        if hdr is None or len(r) < len(hdr): continue
        nm = r[hdr['Kernel Name']].split('(')[0][-40:]
        v = float(r[hdr['Metric Value']].replace(',', '')); u = r[hdr['Metric Unit']]
        v = v / 1000 if u == 'ns' else v * 1000 if u == 'ms' else v
        agg.setdefault(nm, []).append(round(v, 1))
    print(f.split('launches_')[1])
    for nm, vs in agg.items(): print('   ', nm, vs)
PY
