#!/bin/bash
# Round-2 final lines on 8 GPUs of one box: the headline at N=8 and BASELINE config 4 (16,384 instances: Merlin V12, LFA V10, radial 9,
# EJ25) at the scripts' own simulation frequencies, by engine type and with every type on every GPU.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
E=merlin_v12,lfa_v10,radial_9,subaru_ej25
$TR --master-port 29541 bench.py --gpus 8 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_r02f_n8.json 2> gpurun_out/n8_r02f.err
$TR --master-port 29542 bench.py --gpus 8 --engine $E --instances 2048 --sim-freq 0 --steps 2 --no-cpu-baseline > gpurun_out/bench_r02f_config4_bytype_ownfreq_n8.json 2>> gpurun_out/n8_r02f.err
$TR --master-port 29544 bench.py --gpus 8 --engine $E --instances 2048 --sim-freq 0 --steps 2 --placement mixed --no-cpu-baseline > gpurun_out/bench_r02f_config4_mixed_ownfreq_n8.json 2>> gpurun_out/n8_r02f.err
for f in gpurun_out/bench_r02f_n8.json gpurun_out/bench_r02f_config4_*_n8.json; do python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print(sys.argv[1], round(d["value"],1), round(d["ms_per_step"],1), [round(x) for x in d.get("per_gpu_busy_ms_per_step",[])])
except Exception as e: print(sys.argv[1], "FAILED", e)
PY
done
tail -5 gpurun_out/n8_r02f.err
