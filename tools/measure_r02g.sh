#!/bin/bash
# phase clocks of the instrumented build, then parity + bench of the variants given on the command line
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
ESB_LIBRARY=$PWD/engine-sim-cli_b200/libenginesim_b200_clk.so timeout 300 python tools/phase_clocks.py 4096 | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print(d['n'], round(d['cycles_per_step']))
    for k,v in d['phases'].items(): print('   %-32s %8d  x%-5.0f = %6d  %.3f'%(k,v['cycles_per_step'],v['marks_per_step'],v['cycles_per_mark'],v['share']))
"
bash tools/measure_r02e.sh "$@"
