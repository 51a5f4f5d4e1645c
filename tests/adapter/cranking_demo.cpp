// cranking_demo.cpp — the reference's OWN start/stop state machine (engine-sim-bridge/src/simulation/CrankingController.cpp,
// compiled from where it lies under /root/reference, never copied) driving the GPU engine through the reference's
// ICombustionEngine interface (include/GpuSimulator.h), in the order SimulationLoop uses (SimulationLoop.cpp:142-157
// applyCrankingDecision, :166-175 throttle / ignition, BridgeSimulator::update).  tests/test_gpu_adapter.py compares the
// phase / RPM trace it prints with engine_sim_cli_b200/cranking.py's vectorised controller on the same engine and seed.
//   usage: cranking_demo <params.bin> <ir.s16> <ticks> <throttle> <ignition 0|1>
#include "GpuSimulator.h"
#include "simulation/CrankingController.h"

#include <cstdio>
#include <cstdlib>
#include <vector>

int main(int argc, char **argv) {
    if (argc < 6) return 2;
    static EsbEngineParams P;
    FILE *f = std::fopen(argv[1], "rb");
    if (!f || std::fread(&P, sizeof(P), 1, f) != 1) { std::fprintf(stderr, "cannot read %s\n", argv[1]); return 1; }
    std::fclose(f);
    std::vector<int16_t> ir;
    f = std::fopen(argv[2], "rb");
    if (!f) return 1;
    int16_t v;
    while (std::fread(&v, 2, 1, f) == 1) ir.push_back(v);
    std::fclose(f);
    const int ticks = std::atoi(argv[3]);
    const double throttle = std::atof(argv[4]);
    const bool ignition = std::atoi(argv[5]) != 0;

    GpuSimulator gpu;
    ICombustionEngine *engine = &gpu;
    ISimulatorConfig cfg;
    cfg.simulationFrequency = 10000;
    if (!engine->create(cfg, nullptr, nullptr)) { std::fprintf(stderr, "create failed: %s\n", engine->getLastError().c_str()); return 3; }
    if (!gpu.loadParams(&P, ir.data(), (int32_t)ir.size(), 0.001f)) { std::fprintf(stderr, "load failed: %s\n", engine->getLastError().c_str()); return 4; }
    CrankingController controller;
    std::printf("[");
    for (int t = 0; t < ticks; ++t) {
        const bool starterButton = (t == 0);   // ReplayTelemetryProvider: one starter pulse on the first frame
        engine->applyTransition(controller.engageStarter(*engine, starterButton, ignition));
        const TransitionDecision d = controller.step(*engine, throttle, ignition);
        engine->applyTransition(d);
        engine->setThrottle(d.effectiveThrottle);
        engine->setIgnition(ignition);
        engine->update(1.0 / 60.0);
        const EngineSimStats st = engine->getStats();
        std::printf("%s[%d, %.17g, %.17g]", t ? ", " : "", (int)engine->getEnginePhase(), st.currentRPM, st.exhaustFlow);
    }
    std::printf("]\n");
    engine->destroy();
    return 0;
}
