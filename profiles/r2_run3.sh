O=gpurun_out; mkdir -p $O
python profiles/prof_norm.py
ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,launch__registers_per_thread,sm__warps_active.avg.pct_of_peak_sustained_active --clock-control none -s 10 -c 11 --csv --log-file $O/r2_norm_launches.csv python profiles/prof_norm.py > /dev/null 2>&1
python - <<'PY'
import csv
rows=list(csv.reader(open('gpurun_out/r2_norm_launches.csv')))
hdr=[i for i,r in enumerate(rows) if r and r[0]=='ID'][0]
for r in rows[hdr+1:]:
    print(r[0], r[4][:60], r[-3], r[-1])
PY
ncu --set full --import-source on --clock-control none -k regex:norm_sentence_tiles -s 2 -c 2 -f -o $O/r2_norm_sentence python profiles/prof_norm.py > /dev/null 2>&1
ls -la $O
