#!/bin/bash
# A/B of library variants: parity subset + headline bench for each library given on the command line (ESB_LIBRARY).
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for lib in "$@"; do
  tag=$(basename $lib .so | sed 's/libenginesim_b200//; s/^_//'); tag=${tag:-main}
  export ESB_LIBRARY=$PWD/engine-sim-cli_b200/$lib
  echo "== $lib"
  timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -3 | tee gpurun_out/pytest_ab_$tag.log
  timeout 300 python bench.py --no-cpu-baseline > gpurun_out/bench_ab_$tag.json 2> gpurun_out/bench_ab_$tag.err
  python -c "import json,sys; l=json.loads(open('gpurun_out/bench_ab_$tag.json').read().strip().splitlines()[-1]); print('$tag', round(l['value'],1), round(l['ms_per_step'],2), 'e2e', round(l['e2e']['value'],1))"
done
