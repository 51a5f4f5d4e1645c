#!/bin/bash
# Round-2 lines on 8 GPUs of one box: the headline at N=8, BASELINE config 4 (16,384 instances: Merlin V12, LFA V10, radial 9, EJ25)
# by engine type and with every type on every GPU, at the scripts' own simulation frequencies and at a forced 10 kHz; the product's
# multi-device render (render.py --devices).
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
E=merlin_v12,lfa_v10,radial_9,subaru_ej25
$TR --master-port 29541 bench.py --gpus 8 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_r02_n8.json 2> gpurun_out/n8.err
$TR --master-port 29542 bench.py --gpus 8 --engine $E --instances 2048 --sim-freq 0 --steps 2 --no-cpu-baseline > gpurun_out/bench_r02_config4_bytype_ownfreq_n8.json 2>> gpurun_out/n8.err
$TR --master-port 29543 bench.py --gpus 8 --engine $E --instances 2048 --sim-freq 10000 --steps 2 --no-cpu-baseline > gpurun_out/bench_r02_config4_bytype_10khz_n8.json 2>> gpurun_out/n8.err
$TR --master-port 29544 bench.py --gpus 8 --engine $E --instances 2048 --sim-freq 0 --steps 2 --placement mixed > gpurun_out/bench_r02_config4_mixed_ownfreq_n8.json 2>> gpurun_out/n8.err
python -m engine_sim_cli_b200.render --script tests/golden/gm_ls.params.bin --ir tests/golden/gm_ls.ir.npy --rpm 1000:6000:64 --duration 1 --devices 0-7 --output /tmp/out8 --telemetry gpurun_out/render_devices_r02.csv > gpurun_out/render_devices_r02.log 2>&1
ls /tmp/out8 | wc -l >> gpurun_out/render_devices_r02.log
for f in gpurun_out/bench_r02_n8.json gpurun_out/bench_r02_config4_*_n8.json; do python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print(sys.argv[1], round(d["value"],1), round(d["ms_per_step"],1), [round(x) for x in d.get("per_gpu_busy_ms_per_step",[])])
except Exception as e: print(sys.argv[1], "FAILED", e)
PY
done
tail -3 gpurun_out/render_devices_r02.log; tail -5 gpurun_out/n8.err
