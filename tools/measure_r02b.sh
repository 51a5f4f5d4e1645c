#!/bin/bash
# Round-2 secondary lines on ONE GPU: configs 2 and 3 (long IR), config 5 with 1 and 16 impulse responses, the state-machine path.
run() { name=$1; shift; python bench.py --steps 3 --warmup 3 --no-cpu-baseline "$@" > gpurun_out/bench_r02_${name}.json 2> gpurun_out/bench_r02_${name}.err; python -c "import json,sys; d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print(sys.argv[2], round(d['value'],1), round(d['ms_per_step'],1))" gpurun_out/bench_r02_${name}.json $name; }
run config2_256 --instances 256
run config3_longir --ir-taps 33705
python bench_fir.py --irs 1 > gpurun_out/bench_fir_config5_r02.json 2> gpurun_out/fir.err; tail -c 400 gpurun_out/bench_fir_config5_r02.json
python bench_fir.py --irs 16 > gpurun_out/bench_fir_config5_16irs_r02.json 2>> gpurun_out/fir.err; tail -c 400 gpurun_out/bench_fir_config5_16irs_r02.json
python tools/bench_session.py 4096 2.0 > gpurun_out/bench_session_r02.jsonl 2> gpurun_out/session.err; cat gpurun_out/bench_session_r02.jsonl
python bench.py > gpurun_out/bench_r02_n1_traffic.json 2>/dev/null; python -c "import json; d=json.load(open('gpurun_out/bench_r02_n1_traffic.json')); print(d['value'], d['roofline']['frac'], d['roofline']['traffic'])"
