#!/bin/bash
# Factorisation at scale: the 1.08 M-tet cuboid (4 degree twist, 3 + 1 ANM
# iterations = 4 factorisations) through gpu_fea with per-kernel-type device
# times.  gpurun --timeout 1200 -- 'bash tools/c61_factor.sh <tag>'
R=$PWD
tag=${1:-c61}
mkdir -p /tmp/c61 && cd /tmp/c61
SANM_LU_PROF=1 timeout 900 $R/cxx/_build/gpu_fea $R/oracle/_ref/config/sys.json \
    $R/bench_configs/cuboid61_twist_neo_incomp.json $R/bench_configs/override_rotate4.json \
    > $R/gpurun_out/$tag.txt 2>&1
grep -E "solving|timing|sparse_prep|lu_wide|lu_asm|lu_gemm|sparse_solve " $R/gpurun_out/$tag.txt | tail -14
