for B in 1 2 3 4; do for C in 512 4096; do for M in fused_graph unfused_graph; do
  python bench.py --workload cfg4 --batch $B --context $C --steps 32 --warmup 4 --no-cpu-baseline --mode $M 2>&1 | grep "^{" | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('B=$B ctx=$C $M', d['ms_per_step'], d['value'])"
done; done; done
