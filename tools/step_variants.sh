#!/bin/bash
# Step-kernel occupancy variants (SANM_STEP_MINB) on the 2.0 M-tet cuboid, one
# bench line per variant.  Run under gpurun:
#   gpurun --timeout 1500 -- 'bash tools/step_variants.sh'
out=gpurun_out/step_variants.txt
: > $out
for cfg in "neohookean_c 5" "neohookean_i 4" "neohookean_i 5" "arap 4" "arap 5"; do
    set -- $cfg
    line=$(SANM_STEP_MINB=$2 python bench.py --steps 8 --warmup 3 --energy $1 \
        --nx 201 --ny 51 --nz 41 --no-solves --no-cpu --no-energies 2>/dev/null | tail -1)
    echo "$1 MINB=$2 $(echo "$line" | python -c '
import json,sys
d=json.loads(sys.stdin.read())
r=d["roofline"]
print("value=%.4g frac=%.4f launch_ms=%s clocks=%s" % (d["value"], r["frac"], r.get("launch_ms_avg"), d.get("clocks",{}).get("sm_mhz")))')" >> $out
done
cat $out
