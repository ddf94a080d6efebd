CMD="python bench.py --workload horseshoe --chains 256 --steps 1 --warmup 1 --adapt 1 --transitions 2 --fixed-step 0.002 --max-tree-depth 8"
timeout 600 $CMD 2>&1 | tail -1 | cut -c1-300 && ncu --metrics gpu__time_duration.sum --clock-control none -s 1500 -c 300 --csv --log-file gpurun_out/launches_hs.csv $CMD > gpurun_out/ncu_hs.log 2>&1
python - <<EOF
import csv,collections
rows=list(csv.reader(l for l in open("gpurun_out/launches_hs.csv") if l.startswith("\"")))
h=rows[0]; ki=h.index("Kernel Name"); vi=h.index("Metric Value")
agg=collections.OrderedDict()
for r in rows[1:]:
    try: v=float(r[vi].replace(",",""))
    except: continue
    a=agg.setdefault(r[ki][:60],[0,0.0]); a[0]+=1; a[1]+=v
for k,a in agg.items(): print(k, a[0], round(a[1]/a[0]/1000,1), "us avg")
EOF
