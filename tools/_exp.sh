set -x
timeout 1500 python -m pytest tests/test_gpu_parity.py tests/test_gpu_baseline_shapes.py tests/test_gpu_multichain.py tests/test_refpin.py tests/test_gpu_ecs.py tests/test_gpu_posterior.py -m gpu -x -q 2>&1 | tail -15 > gpurun_out/exp5_tests.log
B="--no-out-of-l2 --no-multi-chain --no-many-chain --no-cpu --steps 6 --warmup 3"
timeout 300 python bench.py $B > gpurun_out/exp5_twolane.json 2> gpurun_out/exp5_twolane.err
timeout 300 python bench.py --workload ecs --steps 5 --warmup 2 --transitions 200 --adapt 500 > gpurun_out/r02_bench_ecs.json 2> gpurun_out/r02_bench_ecs.err
B200MC_ECS_ROWS_SMEM=0 timeout 300 python bench.py --workload ecs --steps 5 --warmup 2 --transitions 200 --adapt 500 --no-cpu > gpurun_out/exp5_ecs_norows.json 2> /dev/null
timeout 300 python bench.py --workload ecs --chains 256 --steps 3 --warmup 1 --transitions 200 --adapt 500 --no-cpu > gpurun_out/r02_bench_ecs_256.json 2> gpurun_out/r02_bench_ecs_256.err
ls -la gpurun_out | tail -6
