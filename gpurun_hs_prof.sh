CMD="python bench.py --workload horseshoe --chains 256 --steps 1 --warmup 1 --adapt 1 --transitions 2 --fixed-step 0.0005 --max-tree-depth 5"
timeout 600 $CMD > gpurun_out/hs_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:gemm3x -s 40 -c 2 -o gpurun_out/prof_gemm $CMD > gpurun_out/ncu_gemm.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:batch_matvec -s 60 -c 1 -o gpurun_out/prof_matvec $CMD > gpurun_out/ncu_mv.log 2>&1
tail -c 300 gpurun_out/hs_plain.log
