#!/bin/bash
# Round-2 final measurement pass on ONE GPU (final kernel of round 2: whole-warp shuffles): headline bench line + reference arm, launch
# list, ncu --set full of the physics kernel at the bench's launch shape, per-engine lines, configs 2 and 3, the state-machine path.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
ncu --set full --clock-control none --import-source on -k regex:esb_physics -s 2 -c 1 -o gpurun_out/prof_physics_r02f python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_phys_r02f.log 2>&1
python tools/ncu_summary.py gpurun_out/prof_physics_r02f.ncu-rep profiles/physics_ncu_r02f.json 500 4096 > /dev/null && cp profiles/physics_ncu_r02f.json gpurun_out/
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_r02f.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_list_r02f.log 2>&1
python bench.py > gpurun_out/bench_r02f_n1.json 2> gpurun_out/bench_r02f_n1.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_r02f_reference.json 2>> gpurun_out/bench_r02f_n1.err
cut -c1-330 gpurun_out/bench_r02f_n1.json; cut -c1-200 gpurun_out/bench_r02f_reference.json
ls -la gpurun_out/*.ncu-rep
bash tools/bench_engines_quick.sh r02f
run() { name=$1; shift; python bench.py --steps 3 --warmup 3 --no-cpu-baseline "$@" > gpurun_out/bench_r02f_${name}.json 2> gpurun_out/bench_r02f_${name}.err; python -c "import json,sys; d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print(sys.argv[2], round(d['value'],1), round(d['ms_per_step'],1))" gpurun_out/bench_r02f_${name}.json $name; }
run config2_256 --instances 256
run config3_longir --ir-taps 33705
python tools/bench_session.py 4096 2.0 > gpurun_out/bench_session_r02f.jsonl 2> gpurun_out/session_r02f.err; cat gpurun_out/bench_session_r02f.jsonl
