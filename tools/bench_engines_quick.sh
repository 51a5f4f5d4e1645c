#!/bin/bash
# Developer tool: one short bench line per engine type (4096 instances, one GPU), at a forced 10 kHz and at the engine's own
# simulation frequency.  Output: gpurun_out/engines_<tag>.jsonl
tag=${1:-r02}
out=gpurun_out/engines_$tag.jsonl
: > $out
for e in gm_ls ferrari_f136 subaru_ej25 lfa_v10 merlin_v12 radial_9 hayabusa_i4 honda_trx520; do
  for f in 10000 0; do
    [ "$e" = gm_ls ] && [ "$f" = 0 ] && continue
    python bench.py --no-cpu-baseline --engine $e --sim-freq $f --steps 2 --warmup 3 2>/dev/null >> $out
  done
done
python - <<PY
import json
for l in open("$out"):
    d=json.loads(l)
    print(d["metric"][55:-1], d["config"]["sim_freq_rank0"], round(d["value"],1), round(d["e2e"]["value"],1), {k:round(v,1) for k,v in d["kernel_ms"].items()})
PY
