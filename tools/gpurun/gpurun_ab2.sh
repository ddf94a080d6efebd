run() { timeout 300 python bench.py --no-cpu "${@:2}" 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$1', round(d['value']), round(d['ms_per_step'],2), d['mean_tree_size'], d['adapt_seconds'])"; }
run sv_base --workload sv --adapt 100 --steps 2 --warmup 1 --transitions 20
B200MC_LIB=$PWD/numpyro_b200/libb200mc_unroll.so run sv_unroll --workload sv --adapt 100 --steps 2 --warmup 1 --transitions 20
run es_base --workload eight_schools --steps 5 --warmup 3
B200MC_LIB=$PWD/numpyro_b200/libb200mc_unroll.so run es_unroll --workload eight_schools --steps 5 --warmup 3
