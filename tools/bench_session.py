"""Developer tool: throughput of the start / stop state machine path for a batch of cold starts, device-side session
(one render call) against the host-paced 60 Hz tick loop (set_controls -> render -> telemetry read every tick).
    python tools/bench_session.py [instances] [seconds]"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import bench
import engine_sim_cli_b200 as esb
from engine_sim_cli_b200 import cranking as ck


def run(n, seconds, on_device):
    blob, ir = bench.load_workload()
    b = esb.BatchSimulator(n, device=0, sim_freq=10000, seed=0)
    b.load_params(blob)
    for c in range(2):
        b.set_impulse_response(c, ir, 0.001)
    trace = {"time": np.array([0.0, 1.0, 2.0]), "throttle": np.array([0.05, 0.4, 0.9]), "rpm": np.array([-1.0, -1.0, -1.0])}
    s = ck.BatchSession(b, sim_freq=10000)
    b.state()
    t0 = time.perf_counter()
    pcm, phases, rpms = (s.run_replay_on_device if on_device else s.run_replay)(trace, seconds=seconds, want_pcm=False)
    b.state()
    dt = time.perf_counter() - t0
    return {"mode": "device-side session (one render call)" if on_device else "host-paced ticks", "instances": n, "engine_seconds_each": seconds,
            "wall_s": dt, "engine_s_per_s": n * seconds / dt, "running_at_end": int((phases[-1] == ck.EnginePhase.Running).sum())}


if __name__ == "__main__":
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
    sec = float(sys.argv[2]) if len(sys.argv) > 2 else 3.0
    for mode in (True, False):
        print(json.dumps(run(n, sec, mode)), flush=True)
