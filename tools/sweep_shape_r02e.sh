#!/bin/bash
# Block size / steps-per-launch sweep of the headline configuration with the final r02e kernel.
cd "$(dirname "$0")/.."
run() { python bench.py --steps 3 --warmup 3 --no-cpu-baseline "$@" 2>/dev/null | python -c "import json,sys; l=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(sys.argv[1:], round(l['value'],1), round(l['ms_per_step'],2))" "$@"; }
for bt in 64 128 224 256; do run --block-threads $bt; done
for spl in 250 1000; do run --steps-per-launch $spl; done
