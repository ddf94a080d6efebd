set -x
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29551 bench.py --gpus 4 --steps 6 --warmup 3 > gpurun_out/r02_bench_rows4.json 2> gpurun_out/r02_bench_rows4.err
