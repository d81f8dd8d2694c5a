L=dvs_of_training_framework_b200/libdvsof.so
for v in P Q; do
  cp _lib_$v.so $L
  echo "variant $v"
  DVSOF_ENCODE_PATH=b python scripts/exp_encoder.py cfg3 2>&1 | tail -1
  ncu --metrics gpu__time_duration.sum --clock-control none -k regex:encode_ -c 40 --csv --log-file gpurun_out/launches_$v.csv env DVSOF_ENCODE_PATH=b python scripts/exp_encoder.py cfg3 > /dev/null 2>&1
  python - <<P
import csv,collections
rows=list(csv.reader(l for l in open('gpurun_out/launches_$v.csv') if l.startswith('"')))
h=rows[0]; ki=h.index('Kernel Name'); vi=h.index('Metric Value')
d=collections.defaultdict(list)
for r in rows[1:]: d[r[ki][:40]].append(float(r[vi].replace(',','')))
for k,v in d.items(): print('   ',k, len(v), sorted(v)[len(v)//2])
P
done
