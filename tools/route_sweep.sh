for B in 2 3 4; do for C in 512 4096; do
 for cfg in "persistent 5" "unfused_graph 5" "unfused_graph 2"; do set -- $cfg
  B200_GEMM_MIN_TOKENS=$2 python bench.py --workload cfg4 --batch $B --context $C --steps 32 --warmup 4 --no-cpu-baseline --mode $1 2>&1 | grep "^{" | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('B=$B ctx=$C $1 gemm_min=$2', d['ms_per_step'], d['value'])"
 done; done; done
