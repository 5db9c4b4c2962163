# bench (both arms) + ncu launch list of the same command
O=gpurun_out; mkdir -p $O
timeout 600 python bench.py > $O/r2_bench.json 2> $O/r2_bench.err; tail -c 2500 $O/r2_bench.json; tail -3 $O/r2_bench.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r2_launches.csv python bench.py --steps 8 --warmup 3 > $O/ncu.log 2>&1
python - <<'PY'
import csv, collections
rows=list(csv.reader(open('gpurun_out/r2_launches.csv')))
h=[i for i,r in enumerate(rows) if r and r[0]=='ID'][0]
agg=collections.defaultdict(lambda:[0,0.0])
for r in rows[h+1:]:
    if len(r)<5: continue
    k=r[4][:70]; agg[k][0]+=1; agg[k][1]+=float(r[-1])
for k,(n,t) in sorted(agg.items(), key=lambda kv:-kv[1][1])[:12]: print("%4d %10.1f us  %8.2f us/launch  %s"%(n,t/1e3,t/1e3/n,k))
PY
