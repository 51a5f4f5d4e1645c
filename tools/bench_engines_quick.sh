# quick throughput check of three engine shapes (8, 16 and 4 lanes per instance): bash tools/bench_engines_quick.sh
for e in gm_ls merlin_v12 lfa_v10 subaru_ej25; do
  python bench.py --steps 3 --warmup 3 --no-cpu-baseline --engine $e 2>/dev/null | E=$e python -c "import json,sys,os; l=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(os.environ['E'], round(l['value'],1), round(l['ms_per_step'],1))"
done
