# Developer sweep of the physics block size with the default workload (run on the GPU box: bash tools/sweep_block_threads.sh 224 256 128)
mkdir -p gpurun_out; : > gpurun_out/sweep_bt.log
for bt in "$@"; do
  python bench.py --steps 3 --warmup 3 --no-cpu-baseline --block-threads "$bt" 2>/dev/null | BT="$bt" python -c "import json,sys,os; l=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(os.environ['BT'], l['value'], l['kernel_ms']['physics'])" >> gpurun_out/sweep_bt.log
done
cat gpurun_out/sweep_bt.log
