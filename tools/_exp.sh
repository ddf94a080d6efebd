timeout 300 python bench.py --workload ecs --steps 5 --warmup 2 --transitions 200 --adapt 500 > gpurun_out/r02_bench_ecs.json 2> gpurun_out/r02_bench_ecs.err
timeout 300 python bench.py --workload ecs --chains 256 --steps 3 --warmup 1 --transitions 200 --adapt 500 --no-cpu > gpurun_out/r02_bench_ecs_256.json 2> gpurun_out/r02_bench_ecs_256.err
