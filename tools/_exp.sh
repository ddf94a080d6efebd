set -x
B="--no-out-of-l2 --no-multi-chain --no-many-chain --no-cpu --steps 6 --warmup 3"
for v in w12p1 w14p1 w12p0; do B200MC_LIB=$PWD/numpyro_b200/lib$v.so timeout 300 python bench.py $B > gpurun_out/exp11_$v.json 2> gpurun_out/exp11_$v.err; done
