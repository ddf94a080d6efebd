for i in 1 2; do
B200MC_LIB=$PWD/gpurun_ab_old.so timeout 200 python bench.py --steps 5 --warmup 3 --no-cpu 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('OLD', d['value'], d['roofline']['us_per_grad_eval'], d['cta0_cycles_per_grad_eval'], d['mean_tree_size'])"
timeout 200 python bench.py --steps 5 --warmup 3 --no-cpu 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('NEW', d['value'], d['roofline']['us_per_grad_eval'], d['cta0_cycles_per_grad_eval'], d['mean_tree_size'])"
done
