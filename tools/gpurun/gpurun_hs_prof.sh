CMD="python bench.py --workload logistic_many --chains 256 --steps 1 --warmup 1 --adapt 20 --transitions 3"
timeout 600 $CMD > gpurun_out/lm_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 600 -c 300 --csv --log-file gpurun_out/launches_lm.csv $CMD > gpurun_out/ncu_lm.log 2>&1
python - <<EOF
import csv,collections
rows=list(csv.reader(l for l in open("gpurun_out/launches_lm.csv") if l.startswith("\"")))
h=rows[0]; ki=h.index("Kernel Name"); vi=h.index("Metric Value")
agg=collections.OrderedDict()
for r in rows[1:]:
    try: v=float(r[vi].replace(",",""))
    except: continue
    a=agg.setdefault(r[ki][:60],[0,0.0]); a[0]+=1; a[1]+=v
for k,a in agg.items(): print(k, a[0], round(a[1]/a[0]/1000,1), "us avg")
EOF
