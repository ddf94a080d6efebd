set -x
timeout 900 python tools/ncu_chain.py capture sv > gpurun_out/ncu_chain_sv.log 2>&1
timeout 900 python tools/ncu_traffic.py capture > gpurun_out/ncu_traffic.log 2>&1
rm -f gpurun_out/*.ncu-rep
B="--no-out-of-l2 --no-multi-chain --no-many-chain --no-cpu --steps 2 --warmup 1"
timeout 300 python bench.py $B > gpurun_out/launch_b.log 2>&1 && timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches.csv python bench.py $B > gpurun_out/launch_ncu.log 2>&1
ls -la gpurun_out | tail -8
