#!/bin/bash
# A/B of FIR kernel builds on BASELINE config 5 (bench_fir.py); libraries under build_old/ (developer scratch, not committed)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for lib in "$@"; do
  echo "== $lib"
  ESB_LIBRARY=$PWD/$lib timeout 300 python bench_fir.py --samples 132300 --reps 2 2>&1 | tail -1
done
