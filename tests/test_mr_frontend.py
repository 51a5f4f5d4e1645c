"""The Piranha-free .mr front-end (engine_sim_cli_b200/mr_frontend.py).

* Language features on a self-contained miniature library (tests/golden/mr_mini), so this part runs anywhere.
* With the reference checkout present: the reference's own script <-> preset fixture pairs
  (es/engines/.../07_gm_ls.mr, 01_honda_trx520.mr, 01_subaru_ej25_eh.mr against engine-sim-bridge/test/fixtures/*.preset.json,
  the files its PresetIsomorphismTests.cpp compares) must agree on every numeric field to 1e-9, every other engine script must
  compile, and the resulting documents must load through the CUDA library's preset loader into the committed golden blobs."""
import importlib.util
import json
import math
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference/engine-sim-bridge"
MINI = os.path.join(ROOT, "tests", "golden", "mr_mini")
needs_ref = pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "es")), reason="reference checkout not present")


def _mr():
    spec = importlib.util.spec_from_file_location("esb_mr_frontend", os.path.join(ROOT, "engine-sim-cli_b200", "mr_frontend.py"))
    m = importlib.util.module_from_spec(spec)
    sys.modules["esb_mr_frontend"] = m
    spec.loader.exec_module(m)
    return m


def test_language_features_on_the_mini_library():
    m = _mr()
    it = m.Interpreter([MINI])
    fs = it.load(os.path.join(MINI, "script.mr"))
    main = it.instantiate("main", [], fs)
    inch = 0.0254
    assert main.get("h") == 0 and isinstance(main.get("h"), int)          # int / int
    assert main.get("t") == pytest.approx(1 / 3.0)
    assert main.get("a") == pytest.approx(4.0 * (60.0 / (4.0 * 12.0)))   # defaults reading other inputs, either direction
    assert main.get("b") == pytest.approx(10.0 * (60.0 / (2.0 * 6.0)))   # an explicit input overrides the parameter block
    assert m.deref(main.get("d")) == 6
    assert main.get("s") == "mini ok"
    assert main.get("neg") == pytest.approx(-inch * 2 - 1)
    f = m.deref(main.get("f"))
    assert f["filterRadius"] == 2.0 and f["samples"] == [[0.0, 5.0], [1.0, 10.0]]     # chained actions, kept sorted
    assert m.deref(main.get("pair").get("first")) is f                                 # outputs are evaluated once
    assert m.deref(main.get("k")) == pytest.approx(0.0222954419978396, rel=1e-12)             # gm_ls.preset.json intakes[0].inputFlowK
    lobe = m.deref(main.get("lobe_out"))
    # Honda TRX520's lobe (01_honda_trx520.mr:104-109): 199 samples, k = 8/3, lift(45 cam degrees) = 50 thou
    assert len(lobe["samples"]) == 199 and lobe["filterRadius"] == pytest.approx(0.0124010236325929, rel=1e-12)
    x = 180 * (3.14159265359 / 180.0) / 4
    i = min(range(199), key=lambda j: abs(lobe["samples"][j][0] - x))
    xs, ys = lobe["samples"][i]
    assert ys == pytest.approx(200 * inch / 1000 * (0.5 + 0.5 * math.cos(8.0 / 3.0 * xs)), rel=1e-9)
    assert max(p[1] for p in lobe["samples"]) == pytest.approx(200 * inch / 1000)


def test_errors_are_reported_with_file_and_reason(tmp_path):
    m = _mr()
    bad = tmp_path / "bad.mr"
    bad.write_text('import "engine_sim.mr"\npublic node main { output x: nothing_here; }\n')
    it = m.Interpreter([MINI])
    inst = it.instantiate("main", [], it.load(str(bad)))
    with pytest.raises(m.MrError, match="unknown name 'nothing_here'"):
        inst.get("x")
    (tmp_path / "syntax.mr").write_text("public node main { input a [float: 1; }")
    with pytest.raises(m.MrError, match="syntax.mr:1"):
        m.Interpreter([MINI]).load(str(tmp_path / "syntax.mr"))
    (tmp_path / "noengine.mr").write_text('import "engine_sim.mr"\npublic node main { output x: 1; }\n')
    with pytest.raises(m.MrError, match="never calls set_engine"):
        m.Interpreter([MINI]).run(str(tmp_path / "noengine.mr"))
    with pytest.raises(m.MrError, match="cannot find import"):
        (tmp_path / "imp.mr").write_text('import "missing_file.mr"\n')
        m.Interpreter([MINI]).load(str(tmp_path / "imp.mr"))


def _compare(a, b, path, out):
    if isinstance(b, dict):
        for k in b:
            if k not in a:
                out.append((path + "/" + k, "missing"))
            else:
                _compare(a[k], b[k], path + "/" + k, out)
    elif isinstance(b, list):
        if len(a) != len(b):
            out.append((path, "length", len(a), len(b)))
        else:
            for i, (x, y) in enumerate(zip(a, b)):
                _compare(x, y, f"{path}[{i}]", out)
    elif isinstance(b, (int, float)) and not isinstance(b, bool):
        if abs(a - b) > 1e-9 * max(abs(b), 1e-12) and abs(a - b) > 1e-14:
            out.append((path, a, b))
    elif a != b:
        out.append((path, a, b))


@needs_ref
@pytest.mark.parametrize("script,fixture", [("engines/atg-video-2/07_gm_ls.mr", "gm_ls"), ("engines/atg-video-1/01_honda_trx520.mr", "01_honda_trx520"),
                                            ("engines/atg-video-2/01_subaru_ej25_eh.mr", "01_subaru_ej25_eh")])
def test_reference_script_preset_pairs(script, fixture):
    m = _mr()
    doc = m.compile_script(os.path.join(REF, "es", script))
    ref = json.load(open(os.path.join(REF, "test", "fixtures", fixture + ".preset.json")))
    diffs = []
    _compare(doc, ref, "", diffs)
    # the serializer writes paths relative to its own working directory; both name the same file under sound-library/
    diffs = [d for d in diffs if not (d[0] == "/sourceScript" or d[0].endswith("impulseResponseFilename"))]
    assert diffs == []
    for got, want in zip(doc["engine"]["exhaustSystems"], ref["engine"]["exhaustSystems"]):
        assert want["impulseResponseFilename"].endswith(got["impulseResponseFilename"])


@needs_ref
def test_every_engine_script_compiles():
    m = _mr()
    import glob
    names = {}
    for p in sorted(glob.glob(os.path.join(REF, "es", "engines", "*", "*.mr"))):
        base = os.path.basename(p)
        if base in ("radial.mr", "M156.mr"):       # libraries without a main node
            continue
        if "kohler" in base:                        # the only script with a speed governor instead of a throttle linkage
            with pytest.raises(m.MrError, match="governor"):
                m.compile_script(p)
            continue
        e = m.compile_script(p)["engine"]
        assert e["cylinderCount"] == sum(b["cylinderCount"] for b in e["cylinderBanks"]) > 0
        assert len(e["ignitionModule"]["firingOrder"]) == e["cylinderCount"]
        names[base] = (e["name"], e["cylinderCount"])
    assert names["08_ferrari_f136_v8.mr"] == ("Ferrari F136", 8) and names["10_lfa_v10.mr"][1] == 10 and names["11_merlin_v12.mr"][1] == 12
    assert len(names) >= 19
    # radial engines: one master rod on the crankshaft, every other rod on a journal of the master (09_radial_9.mr:145-165)
    e = m.compile_script(os.path.join(REF, "es", "engines", "atg-video-2", "09_radial_9.mr"))["engine"]
    rods = [b["cylinders"][0]["connectingRod"] for b in e["cylinderBanks"]]
    assert [r["masterCylinder"] for r in rods] == [-1] + [0] * 8
    assert rods[0]["slaveThrow"] == pytest.approx(3.5 * 0.0254) and rods[0]["length"] == pytest.approx(16 * 0.0254)
    assert all(r["length"] == pytest.approx(12.5 * 0.0254) for r in rods[1:])
    assert [r["slaveJournalAngle"] for r in rods[1:]] == pytest.approx([k / 9.0 * 2 * 3.141592653589793 for k in range(1, 9)])


@needs_ref
def test_compiled_scripts_load_into_the_committed_blobs(tmp_path):
    """script -> front-end -> document -> the library's preset loader (csrc/preset.cpp) reproduces tests/golden/*.params.bin,
    including the per-cylinder collector plumbing of the inline four (cylinders alternate between two exhaust systems)."""
    m = _mr()
    csrc = os.path.join(ROOT, "engine-sim-cli_b200", "csrc")
    exe = str(tmp_path / "preset2bin")
    subprocess.run(["g++", "-O2", "-std=c++17", "-I" + os.path.join(ROOT, "include"), "-I" + csrc, os.path.join(csrc, "preset2bin.cpp"),
                    os.path.join(csrc, "preset.cpp"), os.path.join(csrc, "mr_script.cpp"), os.path.join(csrc, "wav.cpp"), "-o", exe], check=True)
    man = json.load(open(os.path.join(ROOT, "tests", "golden", "mr_manifest.json")))
    for name, info in man.items():
        doc = m.compile_script(os.path.join(REF, info["script"]))
        j = tmp_path / (name + ".json")
        j.write_text(json.dumps(doc))
        out = tmp_path / (name + ".bin")
        subprocess.run([exe, str(j), str(out)], check=True, capture_output=True)
        assert out.read_bytes() == open(os.path.join(ROOT, "tests", "golden", name + ".params.bin"), "rb").read(), name
    hay = m.compile_script(os.path.join(REF, man["hayabusa_i4"]["script"]))["engine"]
    assert [c["exhaustSystemIndex"] for b in hay["cylinderBanks"] for c in b["cylinders"]] == [0, 1, 0, 1]


# ---- the library's own front-end (csrc/mr_script.cpp, what EngineSimLoadScript("*.mr") runs): no Python in the loop ----
def _compile_c(path):
    import engine_sim_cli_b200 as esb
    return esb.compile_script(path)


def test_library_front_end_on_the_mini_library(tmp_path):
    """EsbCompileScript on the self-contained mini library equals the Python twin, and its errors carry file and reason."""
    import engine_sim_cli_b200 as esb
    m = _mr()
    src = os.path.join(MINI, "twin.mr")          # a complete (self-written) engine on the mini library
    doc = _compile_c(src)
    out = []
    _compare(doc, m.compile_script(src), "", out)
    assert out == []
    e = doc["engine"]
    assert (e["name"], e["cylinderCount"], e["exhaustSystemCount"]) == ("Mini twin 650", 2, 1)
    assert e["ignitionModule"]["firingOrder"] == pytest.approx([0.0, 2 * 3.14159265359])
    assert e["exhaustSystems"][0]["impulseResponseFilename"] == "sound-library/test/tiny_ir.wav"
    (tmp_path / "noengine.mr").write_text('import "engine_sim.mr"\npublic node main { output x: 1; }\n')
    import shutil
    shutil.copy(os.path.join(MINI, "engine_sim.mr"), tmp_path / "engine_sim.mr")
    with pytest.raises(esb.EngineSimError, match="never calls set_engine") as e:
        _compile_c(str(tmp_path / "noengine.mr"))
    assert e.value.code == -6                                    # ESIM_ERROR_SCRIPT_COMPILATION
    (tmp_path / "syntax.mr").write_text("public node main { input a [float: 1; }")
    with pytest.raises(esb.EngineSimError, match="syntax.mr:1"):
        _compile_c(str(tmp_path / "syntax.mr"))
    with pytest.raises(esb.EngineSimError, match="no such script"):
        _compile_c(str(tmp_path / "missing.mr"))


@needs_ref
@pytest.mark.parametrize("script,fixture", [("engines/atg-video-2/07_gm_ls.mr", "gm_ls"), ("engines/atg-video-1/01_honda_trx520.mr", "01_honda_trx520"),
                                            ("engines/atg-video-2/01_subaru_ej25_eh.mr", "01_subaru_ej25_eh")])
def test_library_front_end_reproduces_the_reference_pairs(script, fixture):
    """The same pin as the Python twin's: the reference's own script <-> preset fixture pairs, every numeric field to 1e-9."""
    doc = _compile_c(os.path.join(REF, "es", script))
    ref = json.load(open(os.path.join(REF, "test", "fixtures", fixture + ".preset.json")))
    diffs = []
    _compare(doc["engine"], {k: v for k, v in ref["engine"].items() if k not in ("exhaustSystems",)}, "engine", diffs)
    for got, want in zip(doc["engine"]["exhaustSystems"], ref["engine"]["exhaustSystems"]):
        _compare(got, {k: v for k, v in want.items() if k != "impulseResponseFilename"}, "exhaust", diffs)
        assert want["impulseResponseFilename"].endswith(got["impulseResponseFilename"])
    assert diffs == []


@needs_ref
def test_library_front_end_equals_the_python_twin_on_every_script():
    import glob
    import engine_sim_cli_b200 as esb
    m = _mr()
    n = 0
    for p in sorted(glob.glob(os.path.join(REF, "es", "engines", "*", "*.mr"))) + [os.path.join(REF, "es", "main.mr")]:
        if os.path.basename(p) in ("radial.mr", "M156.mr"):
            continue
        try:
            want = m.compile_script(p)
        except m.MrError as e:
            with pytest.raises(esb.EngineSimError, match="governor"):
                _compile_c(p)
            continue
        out = []
        _compare(_compile_c(p), want, "", out)
        assert out == [], (p, out[:4])
        n += 1
    assert n >= 20
