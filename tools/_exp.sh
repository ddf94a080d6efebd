set -x
timeout 900 python tools/ncu_traffic.py capture > gpurun_out/ncu_traffic.log 2>&1
ncu -i gpurun_out/glm_r02.ncu-rep --page source --print-source cuda,sass --csv 2>/dev/null | gzip > gpurun_out/glm_r02_source.csv.gz
rm -f gpurun_out/*.ncu-rep
ls -la gpurun_out | tail -4
