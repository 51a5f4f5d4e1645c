// adapter_demo.cpp — drives include/GpuSimulator.h strictly through the reference's ISimulator /
// ICombustionEngine pointers, the way SimulationLoop + SyncPullStrategy do (SimulationLoop.cpp:142-216,
// SyncPullStrategy.cpp:149-161): per tick set controls, renderOnDemand(735 frames).  Built against the
// reference's own headers (only where /root/reference exists); the binary travels to the GPU box and
// tests/test_gpu_adapter.py checks its output against the oracle.
//   usage: adapter_demo <params.bin> <ir.s16> <ticks> <rpm> <out.f32>
#include "GpuSimulator.h"

#include <cmath>
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
    const double rpm = std::atof(argv[4]);

    GpuSimulator gpu;
    ICombustionEngine *engine = &gpu; // callers only ever see the reference's interfaces
    ISimulator *sim = engine;
    ISimulatorConfig cfg;               // reference defaults (EngineSimTypes.h:97-105)
    cfg.simulationFrequency = 10000;
    if (!sim->create(cfg, nullptr, nullptr)) { std::fprintf(stderr, "create failed: %s\n", sim->getLastError().c_str()); return 3; }
    if (!gpu.loadParams(&P, ir.data(), (int32_t)ir.size(), 0.001f)) { std::fprintf(stderr, "load failed: %s\n", sim->getLastError().c_str()); return 4; }
    engine->setIgnition(true);
    engine->applyTransition(TransitionDecision{EnginePhase::Cranking, true, 0.0, true});
    gpu.configureDyno(true, true, rpm, 5000.0);
    sim->start();
    std::vector<float> all, buf(2 * 735);
    long totalWritten = 0;
    for (int t = 0; t < ticks; ++t) {
        sim->setThrottle(std::fmin(1.0, t / 30.0)); // 0 -> 1 over 0.5 s (SimulationLoop.cpp:34-35,49-50)
        int32_t written = 0;
        if (!sim->renderOnDemand(buf.data(), 735, &written)) { std::fprintf(stderr, "render failed: %s\n", sim->getLastError().c_str()); return 5; }
        totalWritten += written;
        all.insert(all.end(), buf.begin(), buf.end());
    }
    const EngineSimStats st = sim->getStats();
    bool threw = false;
    try { sim->renderOnDemand(nullptr, 735, nullptr); } catch (const std::exception &) { threw = true; }
    f = std::fopen(argv[5], "wb");
    std::fwrite(all.data(), sizeof(float), all.size(), f);
    std::fclose(f);
    std::printf("{\"name\": \"%s\", \"rpm\": %.9f, \"written\": %ld, \"phase\": %d, \"channels\": %d, \"simfreq\": %d, \"null_buffer_throws\": %s}\n",
                sim->getName(), st.currentRPM, totalWritten, (int)sim->getEnginePhase(), st.activeChannels, sim->getSimulationFrequency(), threw ? "true" : "false");
    sim->destroy();
    sim->destroy();
    return 0;
}
