"""Developer tool: physics step time against the number of instances (is the kernel latency- or throughput-bound?)."""
import json
import sys
import physics_breakdown as pb

if __name__ == "__main__":
    for n in [int(x) for x in sys.argv[1:]] or [512, 1024, 2048, 4096]:
        print(json.dumps(pb.run(n, 8, 24, steps=300, reps=2, settle=600)), flush=True)
