#!/bin/bash
# End-of-round measurements on one B200 (gpurun): the driver's bench command,
# the reference arm, the ncu launch list of the bench's step, and the DRAM
# traffic of the order-step kernel at the contract size.
#   gpurun --timeout 2400 -- 'bash tools/final_measure.sh'
O=gpurun_out
SHORT="--steps 2 --warmup 1 --no-solves --no-cpu --no-energies"
python bench.py --steps 20 --warmup 5 > $O/bench_n1.json 2> $O/bench_n1.err
tail -c 600 $O/bench_n1.json
python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_reference.json 2> $O/bench_reference.err
tail -c 400 $O/bench_reference.json
python bench.py $SHORT > $O/bench_short.json 2> $O/bench_short.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
    --log-file $O/launches.csv python bench.py $SHORT > $O/ncu_launches.log 2>&1
python bench.py $SHORT > /dev/null 2>&1 && \
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum \
    --clock-control none --kernel-name-base demangled -k regex:StepKernel -s 19 -c 19 --csv \
    --log-file $O/step_dram.csv python bench.py $SHORT > $O/ncu_step_dram.log 2>&1
python bench.py $SHORT > /dev/null 2>&1 && \
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:StepKernel -s 29 -c 1 -f \
    -o $O/step_full python bench.py $SHORT > $O/ncu_step_full.log 2>&1
ls -la $O | tail -12
