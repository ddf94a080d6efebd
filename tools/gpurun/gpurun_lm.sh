# many-chain logistic: correctness of the persistent residual GEMM, then timing A/B, then launch list
python -m pytest tests/test_gpu_batch_engine.py tests/test_gpu_gemm.py -m gpu -q -x > gpurun_out/lm_tests.log 2>&1; tail -3 gpurun_out/lm_tests.log
B200MC_GEMM_PERSIST=0 timeout 300 python bench.py --workload logistic_many --chains 256 --steps 5 --warmup 3 --no-cpu > gpurun_out/lm_old.json 2> gpurun_out/lm_old.err; tail -c 400 gpurun_out/lm_old.err
timeout 300 python bench.py --workload logistic_many --chains 256 --steps 5 --warmup 3 --no-cpu > gpurun_out/lm_new.json 2> gpurun_out/lm_new.err; tail -c 400 gpurun_out/lm_new.err
python - <<PY
import json
for f in ("lm_old","lm_new"):
    try:
        d=json.loads(open("gpurun_out/%s.json"%f).read().strip().splitlines()[-1]); print(f, round(d["value"]), d["ms_per_step"], d.get("roofline"))
    except Exception as e: print(f, "failed", e)
PY
CMD="python bench.py --workload logistic_many --chains 256 --steps 1 --warmup 1 --adapt 20 --transitions 3 --no-cpu"
ncu --metrics gpu__time_duration.sum --clock-control none -s 600 -c 300 --csv --log-file gpurun_out/launches_lm.csv $CMD > gpurun_out/ncu_lm.log 2>&1
python - <<PY
import csv,collections
rows=list(csv.reader(l for l in open("gpurun_out/launches_lm.csv") if l.startswith("\"")))
h=rows[0]; ki=h.index("Kernel Name"); vi=h.index("Metric Value")
agg=collections.OrderedDict()
for r in rows[1:]:
    try: v=float(r[vi].replace(",",""))
    except: continue
    a=agg.setdefault(r[ki][:60],[0,0.0]); a[0]+=1; a[1]+=v
for k,a in agg.items(): print(k, a[0], round(a[1]/a[0]/1000,1), "us avg")
PY
