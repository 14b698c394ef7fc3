#!/bin/bash
# Back-solve latency experiments on Bob (gpurun): launch chain vs CUDA graph
# (SANM_LU_SOLVE_BENCH), pre-wait hoisting on/off, subtree-collapse settings.
out=gpurun_out/solve_latency.txt
: > $out
run() {
  echo "=== $*" >> $out
  env "$@" SANM_LU_SOLVE_BENCH=1 python tools/solve_profile.py bob.json 2>&1 | grep -E "lu-bench|stats|sparse_solve|sparse_prep|solve_expansion" >> $out
}
run SANM_LU_COLLAPSE_NR=104 SANM_LU_NO_HOIST=1
run SANM_LU_COLLAPSE_NR=104
run SANM_LU_COLLAPSE_NR=0
run SANM_LU_COLLAPSE_NR=160 SANM_LU_SMALL_NR=160 SANM_LU_COLLAPSE_NC=128
cat $out
python -m pytest tests/test_sparse_solver.py -m gpu -x -q 2>&1 | tail -2
