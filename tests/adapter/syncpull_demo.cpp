// syncpull_demo.cpp — SURVEY.md §8(f) rank 4: the GPU engine behind the reference's UNMODIFIED host code.
// Compiled together with, from where they lie under /root/reference/engine-sim-bridge (never copied):
//   src/strategy/SyncPullStrategy.cpp      the lock-step audio strategy (the hardware callback calls render(), which pulls
//                                          ISimulator::renderOnDemand and, on a short read, update(1/sr) + retry)
//   src/simulation/CrankingController.cpp  the start/stop state machine
//   src/hardware/NullAudioHardwareProvider.cpp, src/common/ConsoleLogger.cpp
// SimulationLoop.cpp itself cannot be built here: it includes simulator/BridgeSimulator.h, which includes the absent
// engine-sim submodule (BridgeSimulator.h:16); this file restates only its per-tick ORDER (SimulationLoop.cpp:516-530:
// applyCrankingDecision -> applyVehicleControls -> audioBuffer.updateSimulation / fillBufferFromEngine -> getStats) and
// plays the part of the audio hardware by calling the strategy's render() with 512-frame buffers, as CoreAudio would.
//   usage: syncpull_demo <params.bin> <ir.s16> <seconds> <throttle> <out.f32>
#include "GpuSimulator.h"
#include "hardware/NullAudioHardwareProvider.h"
#include "simulation/CrankingController.h"
#include "strategy/SyncPullStrategy.h"

#include <cstdio>
#include <cstdlib>
#include <vector>

namespace {
struct CountingLogger : ILogging {   // keeps stdout machine-readable; warnings and errors go to stderr and are counted
    int warnings = 0, errors = 0;
    uint32_t mask = LogMask::ALL;
    void setMask(uint32_t m) override { mask = m; }
    uint32_t getMask() const override { return mask; }
protected:
    void _write(uint32_t m, const std::string &msg) override {
        const uint32_t level = m & 0xFFFF0000;
        if (level == LogMask::WARN) ++warnings;
        if (level == LogMask::ERROR) ++errors;
        if (level == LogMask::WARN || level == LogMask::ERROR) std::fprintf(stderr, "%s\n", msg.c_str());
    }
};
}

int main(int argc, char **argv) {
    if (argc < 6) return 2;
    static EsbEngineParams P;
    FILE *f = std::fopen(argv[1], "rb");
    if (!f || std::fread(&P, sizeof(P), 1, f) != 1) return 1;
    std::fclose(f);
    std::vector<int16_t> ir;
    f = std::fopen(argv[2], "rb");
    if (!f) return 1;
    int16_t v;
    while (std::fread(&v, 2, 1, f) == 1) ir.push_back(v);
    std::fclose(f);
    const double seconds = std::atof(argv[3]);
    const double throttle = std::atof(argv[4]);

    GpuSimulator gpu;
    ISimulatorConfig cfg;
    cfg.simulationFrequency = 10000;
    if (!gpu.create(cfg, nullptr, nullptr)) return 3;
    if (!gpu.loadParams(&P, ir.data(), (int32_t)ir.size(), 0.001f)) return 4;
    ISimulator *simulator = &gpu;
    ICombustionEngine *engine = &gpu;

    CountingLogger logger;
    SyncPullStrategy strategy(&logger, nullptr);
    AudioBufferConfig bufferConfig;
    NullAudioHardwareProvider hardware;
    AudioStreamFormat format;
    format.sampleRate = EngineSimDefaults::SAMPLE_RATE;
    if (!hardware.initialize(format) || !strategy.initialize(bufferConfig, format.sampleRate)) return 5;
    strategy.prepareBuffer();
    simulator->start();
    if (!strategy.startPlayback(simulator) || !hardware.startPlayback()) return 6;

    CrankingController controller;
    const int sampleRate = format.sampleRate, hwFrames = 512;
    const long long totalFrames = (long long)(seconds * sampleRate);
    std::vector<float> buffer((size_t)hwFrames * 2), all;
    all.reserve((size_t)totalFrames * 2);
    long long frames = 0, nextTick = 0;
    int tick = 0;
    std::printf("{\"trace\": [");
    while (frames < totalFrames) {
        if (frames >= nextTick) {   // the 60 Hz control tick
            engine->applyTransition(controller.engageStarter(*engine, tick == 0, true));
            const TransitionDecision d = controller.step(*engine, throttle, true);
            engine->applyTransition(d);
            simulator->setThrottle(d.effectiveThrottle);
            engine->setIgnition(true);
            strategy.updateSimulation(simulator, 1000.0 / 60.0);
            strategy.fillBufferFromEngine(simulator, sampleRate / 60);
            const EngineSimStats st = simulator->getStats();
            std::printf("%s[%d, %.17g]", tick ? ", " : "", (int)engine->getEnginePhase(), st.currentRPM);
            ++tick;
            nextTick = (long long)tick * sampleRate / 60;
        }
        AudioBufferView view(buffer.data(), hwFrames, 2);   // the audio hardware's pull
        if (!strategy.render(view)) return 7;
        all.insert(all.end(), buffer.begin(), buffer.end());
        frames += hwFrames;
    }
    const Diagnostics::Snapshot snap = strategy.diagnostics().getSnapshot();
    std::printf("], \"frames\": %lld, \"warnings\": %d, \"errors\": %d, \"lastFramesRendered\": %d, \"lastFramesRequested\": %d}\n",
                frames, logger.warnings, logger.errors, snap.lastFramesRendered, snap.lastFramesRequested);
    strategy.stopPlayback(simulator);
    simulator->stop();
    f = std::fopen(argv[5], "wb");
    if (!f) return 8;
    std::fwrite(all.data(), sizeof(float), all.size(), f);
    std::fclose(f);
    simulator->destroy();
    return 0;
}
