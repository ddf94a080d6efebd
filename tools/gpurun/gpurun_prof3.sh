SEC="--section SpeedOfLight --section WarpStateStats --section SchedulerStats --section Occupancy --section MemoryWorkloadAnalysis --section LaunchStats"
for W in sv eight_schools; do
  CMD="python bench.py --workload $W --steps 1 --warmup 1 --adapt 6 --transitions 2 --no-cpu"
  timeout 200 ncu $SEC --clock-control none -k regex:chain_kernel -s 3 -c 1 -o /tmp/chain_$W -f $CMD > gpurun_out/ncu_p3_$W.log 2>&1; tail -1 gpurun_out/ncu_p3_$W.log
  ncu -i /tmp/chain_$W.ncu-rep --page raw --csv > gpurun_out/chain_${W}_raw.csv 2>/dev/null
done
ls -la gpurun_out/*.csv
