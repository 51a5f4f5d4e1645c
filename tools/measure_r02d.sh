#!/bin/bash
# Round-2 pass d: searched gas schedule + warp-uniform solver sweeps.  Quick parity subset, headline bench, phase clocks, slot-sync variant.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -6 | tee gpurun_out/pytest_gpu_r02d_parity.log
timeout 300 python bench.py --no-cpu-baseline > gpurun_out/bench_r02d_n1.json 2> gpurun_out/bench_r02d_n1.err
cut -c1-400 gpurun_out/bench_r02d_n1.json
ESB_LIBRARY=$PWD/engine-sim-cli_b200/libenginesim_b200_sw.so timeout 300 python bench.py --no-cpu-baseline > gpurun_out/bench_r02d_syncwarp.json 2> gpurun_out/bench_r02d_syncwarp.err
cut -c1-200 gpurun_out/bench_r02d_syncwarp.json
ESB_LIBRARY=$PWD/engine-sim-cli_b200/libenginesim_b200_clk.so timeout 300 python tools/phase_clocks.py 4096 256 > gpurun_out/phase_clocks_r02d.json 2> gpurun_out/phase_clocks_r02d.err
cat gpurun_out/phase_clocks_r02d.json
