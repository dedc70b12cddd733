for lib in "" _pb_b3 _pb_b2 _pb_b5 _pb_f8; do
  GSR_B200_LIB=$PWD/gaussian-splatting-npu_b200/gsr_b200/libgsr_b200$lib.so ncu --metrics gpu__time_duration.sum --clock-control none -k regex:batch --csv --log-file gpurun_out/bk$lib.csv python tools/batch_probe.py c2 8 > /dev/null 2>&1
  python - <<PY
import csv,collections
rows=list(csv.DictReader(l for l in open("gpurun_out/bk$lib.csv") if l.startswith('"')))
agg=collections.defaultdict(list)
for r in rows: agg[r['Kernel Name'][:40]].append(float(r['Metric Value'])/1e3)
print("$lib", {k: round(sorted(v)[len(v)//2],1) for k,v in agg.items()})
PY
done
