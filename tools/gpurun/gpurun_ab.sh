run() { timeout 200 python bench.py --steps 5 --warmup 3 --no-cpu 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$1', round(d['value']), round(d['roofline']['us_per_grad_eval'],2), {k:(round(v) if not isinstance(v,(list,dict)) else v) for k,v in d['cta0_cycles_per_grad_eval'].items()}, d['mean_tree_size'])"; }
B200MC_GLM_PINNED=0 run NEW_p0
B200MC_GLM_PINNED=0 B200MC_GLM_SMEM_PAD=30000 run NEW_p0_pad30k
B200MC_GLM_PINNED=0 B200MC_GLM_SMEM_PAD=60000 run NEW_p0_pad60k
B200MC_GLM_PINNED=0 B200MC_GLM_SMEM_PAD=72000 run NEW_p0_pad72k
B200MC_GLM_PINNED=8 run NEW_p8
B200MC_GLM_PINNED=9 run NEW_p9
