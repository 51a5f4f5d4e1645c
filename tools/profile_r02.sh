#!/bin/bash
# Round-2 measurement pass on the GPU box: bench lines, launch list, ncu --set full captures of the physics and the FIR kernel.
set -x
python bench.py > gpurun_out/bench_r02_n1.json 2> gpurun_out/bench_r02_n1.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_r02_reference.json 2>> gpurun_out/bench_r02_n1.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_r02.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_list.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:esb_physics -s 2 -c 1 -o gpurun_out/prof_physics_r02 python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_phys.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:esb_fir -s 2 -c 1 -o gpurun_out/prof_fir_r02 python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_fir.log 2>&1
ls -la gpurun_out/*.ncu-rep
