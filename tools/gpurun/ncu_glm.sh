python tools/gpurun/ncu_glm.py 2>&1 | tail -2
timeout 500 ncu --set full --clock-control none -k regex:glm_kernel -c 6 -o /tmp/glm_v8 -f python tools/gpurun/ncu_glm.py > gpurun_out/ncu_glm_v8.log 2>&1; tail -3 gpurun_out/ncu_glm_v8.log
ncu -i /tmp/glm_v8.ncu-rep --page raw --csv > gpurun_out/glm_v8_raw.csv 2>/dev/null; wc -c gpurun_out/glm_v8_raw.csv
