#!/bin/bash
# Does the REFERENCE keep its own ANM step counts when only its rounding
# changes?  Builds the unmodified reference (oracle/Makefile) three ways --
#   v3     -march=x86-64-v3  (AVX2 + FMA contraction: Eigen packets of 4)  = oracle/_ref
#   v2     -march=x86-64-v2  (SSE: Eigen packets of 2, no FMA)
#   nofma  -march=x86-64-v3 -ffp-contract=off
# -- identical sources, identical algorithm, identical inputs; only the order /
# fusion of floating-point operations inside Eigen's reductions and products
# differs (relative effect ~1e-16 per operation).  Runs the BASELINE configs
# and prints the per-stage iteration counts of each build.
#   usage: tools/step_count_experiment.sh > profiles/r2_step_count_experiment.txt
set -e
ROOT=$(cd "$(dirname "$0")/.." && pwd)
cd "$ROOT/oracle"
make -s -j"$(nproc)" all >/dev/null
make -s -j"$(nproc)" OUT=/tmp/ref_v2 ARCH=x86-64-v2 /tmp/ref_v2/ref_fea >/dev/null
make -s -j"$(nproc)" OUT=/tmp/ref_nofma ARCH=x86-64-v3 EXTRA=-ffp-contract=off /tmp/ref_nofma/ref_fea >/dev/null
CFG="$ROOT/oracle/_ref/config"
run() {  # build-name binary configs...
    local name=$1 bin=$2; shift 2
    local d; d=$(mktemp -d)
    local args=()
    for c in "$@"; do args+=("$CFG/$c"); done
    (cd "$d" && "$bin" "${args[@]}" > out.txt 2>&1) || echo "  $name: rc=$?"
    local iters; iters=$(grep -o 'iter=[0-9]*' "$d/out.txt" | cut -d= -f2 | tr '\n' ' ')
    local frms; frms=$(grep -o '"force_rms_recomp": *[0-9.e+-]*' "$d"/*.json | tail -1 | sed 's/.*: *//')
    local disp; disp=$(grep -o '"displacement": *[0-9.e+-]*' "$d"/*.json | tail -1 | sed 's/.*: *//')
    printf '  %-6s iterations per stage: [ %s]  displacement %s  force_rms %s\n' "$name" "$iters" "$disp" "$frms"
    rm -rf "$d"
}
for cfg in "bob.json" "armadillo_small.json armadillo_small_bend_override.json" \
           "cuboid_twist.json override_order20.json" "armadillo_small.json override_neo_incomp.json"; do
    echo "config: sys.json $cfg"
    run v3 "$ROOT/oracle/_ref/ref_fea" sys.json $cfg
    run v2 /tmp/ref_v2/ref_fea sys.json $cfg
    run nofma /tmp/ref_nofma/ref_fea sys.json $cfg
done
