set -x
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -6 > gpurun_out/r02_pytest_gpu_v4.log
timeout 300 python bench.py --workload eight_schools --chains 4 --adapt 1000 --transitions 1000 --steps 3 --warmup 1 > gpurun_out/r02_bench_eight_schools_4.json 2> gpurun_out/r02_bench_eight_schools_4.err
timeout 300 python bench.py --workload eight_schools > gpurun_out/r02_bench_eight_schools_16384.json 2> gpurun_out/r02_bench_eight_schools_16384.err
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02_smoke.log 2>&1
