# round-end style validation: GPU tests, smoke, both bench arms, launch list of the default bench
python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; tail -2 gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 600 python bench.py --impl reference > gpurun_out/bench_v8_ref.json 2> gpurun_out/bench_v8_ref.err; tail -c 200 gpurun_out/bench_v8_ref.err
timeout 600 python bench.py > gpurun_out/bench_v8.json 2> gpurun_out/bench_v8.err; tail -c 200 gpurun_out/bench_v8.err
python - <<PY
import json
for f in ("bench_v8","bench_v8_ref"):
    d=json.loads(open("gpurun_out/%s.json"%f).read().strip().splitlines()[-1]); print(f, round(d["value"],1), d["e2e"], (d.get("roofline") or {}).get("frac"), d.get("cpu_baseline",{}).get("value"), d.get("clocks"))
PY
