set -x
B200MC_LIB=$PWD/numpyro_b200/libes.so timeout 200 python bench.py --workload eight_schools --steps 6 --warmup 3 > gpurun_out/exp15_es_only.json 2>/dev/null
timeout 200 python bench.py --workload eight_schools --steps 6 --warmup 3 > gpurun_out/exp15_es_base.json 2>/dev/null
