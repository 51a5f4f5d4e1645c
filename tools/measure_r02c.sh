#!/bin/bash
# Round-2 late pass: whole GPU suite, config 5 (1 and 16 impulse responses) with the rebuilt FIR kernel, one ncu --set full capture of it,
# then the headline bench line.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1700 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 | tee gpurun_out/pytest_gpu_r02c.log
timeout 300 python bench_fir.py > gpurun_out/bench_fir_config5_r02c.json 2> gpurun_out/bench_fir_r02c.err
timeout 300 python bench_fir.py --irs 16 > gpurun_out/bench_fir_config5_16irs_r02c.json 2>> gpurun_out/bench_fir_r02c.err
timeout 600 ncu --set full --clock-control none --import-source on -k regex:esb_fir -s 3 -c 1 -o gpurun_out/prof_fir_r02c python bench_fir.py --channels 4096 --samples 44100 --check-channels 0 > gpurun_out/ncu_fir_r02c.log 2>&1
timeout 600 python bench.py > gpurun_out/bench_r02c_n1.json 2> gpurun_out/bench_r02c_n1.err
tail -c 600 gpurun_out/bench_fir_config5_r02c.json; tail -c 300 gpurun_out/bench_fir_r02c.err; cut -c1-300 gpurun_out/bench_r02c_n1.json
