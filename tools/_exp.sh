set -x
timeout 1200 python -m pytest tests/test_gpu_ecs.py -m gpu -x -q 2>&1 | tail -40 > gpurun_out/exp3_ecs_tests.log
timeout 600 python bench.py --workload eight_schools --chains 4 --adapt 1000 --transitions 1000 --steps 3 --warmup 1 > gpurun_out/r02_bench_eight_schools_4.json 2> gpurun_out/r02_bench_eight_schools_4.err
timeout 900 python bench.py --workload sv > gpurun_out/r02_bench_sv.json 2> gpurun_out/r02_bench_sv.err
timeout 900 python tools/ncu_chain.py capture eight_schools > gpurun_out/ncu_chain_es.log 2>&1
ls -la gpurun_out | tail -20
