# Re-measures the secondary bench lines (BASELINE configs 2 and 3, the .mr engines) on one GPU: bash tools/bench_configs.sh <tag>
tag=${1:-r01c}; mkdir -p gpurun_out
run() { name=$1; shift; python bench.py --steps 3 --warmup 3 --no-cpu-baseline "$@" > gpurun_out/bench_${tag}_${name}.json 2> gpurun_out/bench_${tag}_${name}.err; python -c "import json,sys; d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print(sys.argv[2], round(d['value'],1), round(d['ms_per_step'],1))" gpurun_out/bench_${tag}_${name}.json $name; }
run config2_256 --instances 256
run config3_longir --ir-taps 33705
run ferrari_f136_4096 --engine ferrari_f136
run subaru_ej25_4096 --engine subaru_ej25
run lfa_v10_4096 --engine lfa_v10
run merlin_v12_4096 --engine merlin_v12
