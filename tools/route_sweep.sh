#!/bin/bash
# decode launch order x linear kernel sweep for small fp16 batches (7B shapes): ms per step, tok/s
# B200_GEMV_MMA=0: the scalar FHFMA consumer loop; B200_GEMM_MIN_TOKENS: fewest token rows b200_linear sends to the GEMM
for B in ${BATCHES:-2 3 4}; do for C in ${CONTEXTS:-512 4096}; do
 for cfg in "persistent 1 9" "fused_graph 0 9" "fused_graph 1 9" "unfused_graph 1 9" "unfused_graph 1 3"; do set -- $cfg
  B200_GEMV_MMA=$2 B200_GEMM_MIN_TOKENS=$3 python bench.py --workload cfg4 --batch $B --context $C --steps 32 --warmup 4 --no-cpu-baseline --mode $1 2>&1 | grep "^{" | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('B=$B ctx=$C $1 gemv_mma=$2 gemm_min=$3', d['ms_per_step'], d['value'])"
 done; done; done
