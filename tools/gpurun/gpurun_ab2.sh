run() { timeout 300 python bench.py --no-cpu "${@:2}" 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$1', round(d['value']), round(d['ms_per_step'],2), d['mean_tree_size'], d['adapt_seconds'])"; }
for T in 28 32; do B200MC_END_THRESH=$T run es_thresh$T --workload eight_schools --steps 5 --warmup 3; done
B200MC_END_THRESH=32 run es_thresh32_65k --workload eight_schools --chains 65536 --steps 3 --warmup 2
B200MC_END_THRESH=1 run es_thresh1_65k --workload eight_schools --chains 65536 --steps 3 --warmup 2
