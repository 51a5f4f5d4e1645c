#!/bin/bash
# A/B of one library variant, then the whole GPU suite on it.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
bash tools/measure_r02e.sh "$1"
export ESB_LIBRARY=$PWD/engine-sim-cli_b200/$1
timeout 1700 python -m pytest tests -m gpu -x -q --durations=8 2>&1 | tail -16 | tee gpurun_out/pytest_gpu_full.log
