run() { python bench.py --steps 3 --warmup 3 --no-cpu-baseline "$@" 2>/dev/null | python -c "import json,sys; l=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(sys.argv[1:], round(l['value'],1), round(l['ms_per_step'],1))" "$@"; }
run --instances 8192 --variant 2
run --instances 8192 --variant 1
run --instances 8192 --variant 0
run --instances 4096 --variant 2
