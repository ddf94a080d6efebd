set -x
timeout 900 python -m pytest tests/test_gpu_ecs.py tests/test_reflocscale.py tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -15 > gpurun_out/exp4_tests.log
timeout 600 python bench.py --workload ecs --steps 5 --warmup 2 --transitions 200 --adapt 500 > gpurun_out/r02_bench_ecs.json 2> gpurun_out/r02_bench_ecs.err
B="--no-out-of-l2 --no-multi-chain --no-many-chain --no-cpu --steps 4 --warmup 3"
for v in 1 2 3; do B200MC_LIB=$PWD/numpyro_b200/libdbg$v.so timeout 300 python bench.py $B > gpurun_out/exp4_dbg$v.json 2> gpurun_out/exp4_dbg$v.err; done
ls -la gpurun_out | tail -8
