# Developer sweep of the physics launch shape (run on the GPU box: bash tools/sweep_physics_shape.sh)
run() { echo "== $*" >> gpurun_out/sweep.log; python bench.py --steps 2 --warmup 3 --no-cpu-baseline "$@" 2>>gpurun_out/sweep.err | python -c "import json,sys; l=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(l['value'], l['ms_per_step'], l['kernel_ms'])" >> gpurun_out/sweep.log; }
mkdir -p gpurun_out; rm -f gpurun_out/sweep.log gpurun_out/sweep.err
run --block-threads 32 --variant 0
run --block-threads 64 --variant 0
run --block-threads 128 --variant 0
run --block-threads 256 --variant 0
run --block-threads 128 --variant 1 --inst-per-warp 2
run --block-threads 128 --variant 2 --inst-per-warp 2
run --block-threads 64 --variant 1 --inst-per-warp 2
run --instances 8192 --block-threads 128 --variant 2
run --instances 8192 --block-threads 128 --variant 1
