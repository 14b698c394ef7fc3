#!/bin/bash
# Step-kernel variants (fused.cu StepKernel<M, IO, V>, SANM_STEP_V / SANM_STEP_MINB)
# on the 2.0 M-tet cuboid, one bench line per variant.  Run under gpurun:
#   gpurun --timeout 1500 -- 'bash tools/step_variants.sh'
out=gpurun_out/step_variants.txt
: > $out
for e in neohookean_c neohookean_i arap; do
  for cfg in "0 4" "0 5" "1 4" "1 5" "1 6"; do
    set -- $cfg
    line=$(SANM_STEP_V=$1 SANM_STEP_MINB=$2 python bench.py --steps 8 --warmup 3 --energy $e \
        --nx 201 --ny 51 --nz 41 --no-solves --no-cpu --no-energies 2>/dev/null | tail -1)
    echo "$e V=$1 MINB=$2 $(echo "$line" | python -c '
import json,sys
d=json.loads(sys.stdin.read())
r=d["roofline"]
print("value=%.4g frac=%.4f launch_ms=%s clocks=%s" % (d["value"], r["frac"], r.get("launch_ms_avg"), d.get("clocks",{}).get("sm_mhz")))')" >> $out
  done
done
cat $out
