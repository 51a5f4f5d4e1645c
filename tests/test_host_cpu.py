"""CPU-only tests of the host side: the C-ABI library loads and exports every symbol the header declares,
fails loudly without a GPU (no CPU fallback), the preset loader's error behaviour on the reference's own
malformed fixtures, and the multi-rank sharding logic of bench.py (gloo, world_size 2)."""
import ctypes
import os
import re
import subprocess
import sys

import numpy as np
import pytest

from conftest import GOLDEN, ROOT

REF_FIXTURES = "/root/reference/engine-sim-bridge/test/fixtures"


def _lib_path():
    import engine_sim_cli_b200 as esb
    if not os.path.exists(esb.library_path()):
        subprocess.run(["make", "-s", "-j8", "-C", os.path.join(ROOT, "engine-sim-cli_b200", "csrc")], check=True)
    return esb.library_path()


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "engine_sim_b200.h")).read()
    declared = set(re.findall(r"\b(Esb[A-Z]\w+|EngineSim[A-Z]\w+)\s*\(", hdr))
    declared -= {"EngineSimConfig", "EngineSimStats"}
    assert len(declared) >= 45
    lib = ctypes.CDLL(_lib_path())
    missing = [s for s in sorted(declared) if not hasattr(lib, s)]
    assert not missing, missing


def test_no_cpu_fallback():
    import torch
    import engine_sim_cli_b200 as esb
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(esb.EngineSimError) as e:
        esb.BatchSimulator(4)
    assert "no CUDA device" in str(e.value)
    with pytest.raises(esb.EngineSimError):
        esb.Simulator()


def _preset2bin():
    exe = os.path.join(ROOT, "build", "preset2bin")
    csrc = os.path.join(ROOT, "engine-sim-cli_b200", "csrc")
    os.makedirs(os.path.dirname(exe), exist_ok=True)
    subprocess.run(["g++", "-O2", "-std=c++17", "-I" + os.path.join(ROOT, "include"), "-I" + csrc, os.path.join(csrc, "preset2bin.cpp"),
                    os.path.join(csrc, "preset.cpp"), os.path.join(csrc, "mr_script.cpp"), os.path.join(csrc, "wav.cpp"), "-o", exe], check=True)
    return exe


@pytest.mark.skipif(not os.path.isdir(REF_FIXTURES), reason="reference checkout not present")
def test_preset_loader_on_reference_fixtures(tmp_path):
    """Same fixtures the reference's PresetDeserializer tests use: good ones load to the committed blobs,
    malformed ones are rejected with an error (PresetExceptions.h), never a crash."""
    exe = _preset2bin()
    for name, fixture in (("gm_ls", "gm_ls.preset.json"), ("subaru_ej25", "01_subaru_ej25_eh.preset.json"), ("honda_trx520", "01_honda_trx520.preset.json")):
        out = tmp_path / (name + ".bin")
        subprocess.run([exe, os.path.join(REF_FIXTURES, fixture), str(out)], check=True, capture_output=True)
        assert out.read_bytes() == open(os.path.join(GOLDEN, name + ".params.bin"), "rb").read()
    for bad in ("malformed.preset.json", "empty_object.preset.json", "no_engine_section.preset.json", "missing_crankshafts.preset.json",
                "zero_cylinder_count.preset.json"):
        r = subprocess.run([exe, os.path.join(REF_FIXTURES, bad), str(tmp_path / "bad.bin")], capture_output=True, text=True)
        assert r.returncode == 1 and "error" in r.stderr.lower(), (bad, r.returncode, r.stderr)


def test_instance_grid_sharding():
    """Rank r owns instances [r*n, (r+1)*n) of the 64 x 64 RPM/throttle grid; shards tile the grid exactly."""
    sys.path.insert(0, ROOT)
    import bench
    rpm_all, thr_all = bench.instance_grid(4096, 0)
    assert rpm_all.min() == 800.0 and rpm_all.max() == 7000.0 and abs(thr_all.min() - 0.1) < 1e-12 and thr_all.max() == 1.0
    parts = [bench.instance_grid(1024, r * 1024) for r in range(4)]
    assert np.array_equal(np.concatenate([p[0] for p in parts]), rpm_all)
    assert np.array_equal(np.concatenate([p[1] for p in parts]), thr_all)


_WORKER = r"""
import os, sys, json
import numpy as np, torch, torch.distributed as dist
sys.path.insert(0, sys.argv[1])
import bench
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
n = 8
rpm, thr = bench.instance_grid(n, rank * n)
# every rank "processes" its shard; aggregate = units of all ranks / max time over ranks, no data-path collective
ms = torch.tensor([10.0 + rank], dtype=torch.float64)
dist.all_reduce(ms, op=dist.ReduceOp.MAX)
gathered = [None] * world
dist.all_gather_object(gathered, rpm.tolist())
if rank == 0:
    print(json.dumps({"ms": float(ms.item()), "rpm": sum(gathered, []), "value": world * n * 0.1 / (float(ms.item()) * 1e-3)}))
dist.destroy_process_group()
"""


def test_two_rank_sharding_gloo(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(_WORKER)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                        "--master-port", "29533", str(script), ROOT], capture_output=True, text=True, timeout=240)
    assert r.returncode == 0, r.stderr[-2000:]
    import json
    line = json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][-1])
    sys.path.insert(0, ROOT)
    import bench
    assert line["ms"] == 11.0  # max over ranks
    assert line["rpm"] == bench.instance_grid(16, 0)[0].tolist()
    assert abs(line["value"] - 2 * 8 * 0.1 / 0.011) < 1e-9


def test_reference_arm_rank_nonzero_is_silent():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0"],
                       capture_output=True, text=True, env=env, timeout=120)
    assert r.returncode == 0 and r.stdout.strip() == ""


@pytest.mark.parametrize("engine,lanes,max_slots", [
    ("gm_ls", 8, (74, 23)), ("gm_ls", 16, (74, 14)), ("gm_ls", 1, None), ("merlin_v12", 8, (108, 19)), ("lfa_v10", 8, (90, 15)),
    ("radial_9", 8, (84, 18)), ("subaru_ej25", 8, (46, 23)), ("subaru_ej25", 4, None), ("hayabusa_i4", 8, None), ("honda_trx520", 8, None),
])
def test_gas_schedule_keeps_every_cells_operation_order(engine, lanes, max_slots):
    """The slot schedule of the fluid part of a step (csrc/gas_schedule.cpp, found by search) must be a re-ordering that no cell can
    notice: every operation of the reference's sequence (PistonEngineSimulator::simulateStep_, archive/TA2_FINDINGS.md:194-199) exactly
    once, no two operations of a slot on one cell, each cell's own operations in the reference's order.  The library checks that itself
    (EsbGasScheduleInfo); here for every golden engine and several lane counts, with the slot counts the search is expected to reach."""
    from engine_sim_cli_b200 import api
    _lib_path()
    blob = open(os.path.join(GOLDEN, engine + ".params.bin"), "rb").read()
    info = api.gas_schedule_info(blob, min_lanes=lanes)
    assert info["rc"] == 0 and info["problem"] == "", info
    assert info["lanes"] >= lanes
    if max_slots:
        assert info["exchange_slots"] <= max_slots[0] and info["relax_slots"] <= max_slots[1], info
    # other substep counts go through the same search
    info4 = api.gas_schedule_info(blob, min_lanes=lanes, fluid_steps=4)
    assert info4["rc"] == 0 and info4["problem"] == "", info4
    assert info4["exchange_slots"] < info["exchange_slots"]
