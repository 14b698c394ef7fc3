#!/bin/bash
# Factor / solve kernel-type times of the small BASELINE meshes (gpurun).
out=gpurun_out/lu_prof.txt
: > $out
for cfg in "bob.json" "armadillo_small.json override_neo_incomp.json"; do
  echo "=== $cfg (SANM_LU_PROF=1)" >> $out
  SANM_LU_PROF=1 python tools/solve_profile.py $cfg 2>&1 | grep -E "^lu_|sparse_prep|sparse_solve|solve_expansion|stats" | grep -v "lu_fwd\|lu_bwd" >> $out
  echo "=== $cfg (no per-kernel events)" >> $out
  python tools/solve_profile.py $cfg 2>&1 | grep -E "sparse_prep|sparse_solve|solve_expansion|wall_s" >> $out
done
cat $out
python -m pytest tests/test_sparse_solver.py tests/test_anm_parity.py -m gpu -x -q 2>&1 | grep -E "passed|failed|rror" | tail -3
