# headline bench (regression check after the sampler changes) + ncu --set full of the two GEMMs
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_v5.json 2> gpurun_out/bench_v5.err; tail -c 300 gpurun_out/bench_v5.err
timeout 600 python bench.py --steps 5 --warmup 3 --impl reference > gpurun_out/bench_v5_ref.json 2> gpurun_out/bench_v5_ref.err; tail -c 300 gpurun_out/bench_v5_ref.err
CMD="python bench.py --workload logistic_many --chains 256 --steps 1 --warmup 1 --adapt 20 --transitions 3 --no-cpu"
ncu --set full --clock-control none --import-source on -k regex:gemm3x -s 200 -c 2 -o gpurun_out/gemm_r01v2 -f $CMD > gpurun_out/ncu_full.log 2>&1; tail -3 gpurun_out/ncu_full.log
ls -la gpurun_out/*.ncu-rep
